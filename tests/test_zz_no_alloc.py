"""ipkgpu_best_prefix_noalloc (include/ipkgpu.h): the allocation-free variant of ipkgpu_best_prefix, checked through the
C ABI against the existing entry point.  Collected last on purpose."""
import numpy as np
import pytest

from ipk_b200 import capi as libipkgpu


def _posteriors(rng, n_nodes, W, sigma):
    p = rng.gamma(0.35, size=(n_nodes, W, sigma)) + 1e-7
    p /= p.sum(axis=2, keepdims=True)
    return np.log10(p.astype(np.float32)).astype(np.float32)


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    # the size check comes before any CUDA call: no device needed
    n_nodes, W = 4097, 1024                                  # 4097 * 1024 > 2^22 columns
    assert n_nodes * W > libipkgpu.NOALLOC_MAX_COLUMNS
    with pytest.raises(libipkgpu.IpkGpuError) as e:
        libipkgpu.best_prefix_noalloc(0, 0, 16, n_nodes, W, 4, 16)
    assert e.value.code == -1 and "exceeds the static scratch" in str(e.value)
    with pytest.raises(libipkgpu.IpkGpuError) as e:
        libipkgpu.best_prefix_noalloc(0, 0, 16, 4, 8, 64, 16)
    assert e.value.code == -1 and "sigma" in str(e.value)


@pytest.mark.gpu
def test_noalloc_best_prefix_equals_best_prefix_twice_in_a_row(ctx):
    import torch

    rng = np.random.default_rng(7)
    # second call smaller than the first and with another alphabet: whatever the first left in the shared scratch must not
    # show; third call at the largest size the scratch takes
    cases = [(37, 1601, 4, 2), (11, 333, 20, 5), (2048, 2048, 4, 2)]
    host = [_posteriors(rng, n, W, sigma) for n, W, sigma, _ in cases]
    dev_in = [torch.from_numpy(h).cuda() for h in host]
    dev_out = [torch.full((n, W + 1), float("nan"), dtype=torch.float32, device="cuda") for n, W, _, _ in cases]
    stream = torch.cuda.Stream()
    torch.cuda.synchronize()

    free = []
    for i, (n, W, sigma, _) in enumerate(cases):
        # first call on the default stream, the others on a stream of the caller
        s = 0 if i == 0 else stream.cuda_stream
        libipkgpu.best_prefix_noalloc(0, s, dev_in[i].data_ptr(), n, W, sigma, dev_out[i].data_ptr())
        torch.cuda.synchronize()
        free.append(torch.cuda.mem_get_info()[0])
    # (the first call may load the module; after that the device's free memory must not move)
    assert free[1] == free[2], f"device free memory moved between calls: {free}"

    for i, (n, W, sigma, bits) in enumerate(cases):
        want = ctx.best_prefix(host[i], np.arange(n, dtype=np.uint32), W, sigma, bits, 3)
        got = dev_out[i].cpu().numpy()
        assert got.shape == want.shape
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), f"case {i}: bits differ from ipkgpu_best_prefix"

    # and once more after the context's own work, same inputs, same array: same bits
    n, W, sigma, bits = cases[0]
    dev_out[0].fill_(float("nan"))
    torch.cuda.synchronize()
    libipkgpu.best_prefix_noalloc(0, stream.cuda_stream, dev_in[0].data_ptr(), n, W, sigma, dev_out[0].data_ptr())
    stream.synchronize()
    want = ctx.best_prefix(host[0], np.arange(n, dtype=np.uint32), W, sigma, bits, 3)
    assert np.array_equal(dev_out[0].cpu().numpy().view(np.uint32), want.view(np.uint32))


@pytest.mark.gpu
def test_noalloc_rejects_host_pointers_and_misalignment():
    import torch

    a = torch.zeros(4 * 8 * 4 + 4, dtype=torch.float32, device="cuda")
    out = torch.zeros(4 * 9, dtype=torch.float32, device="cuda")
    hostbuf = np.zeros(4 * 8 * 20, np.float32)
    with pytest.raises(libipkgpu.IpkGpuError) as e:
        libipkgpu.best_prefix_noalloc(0, 0, hostbuf.ctypes.data, 4, 8, 20, out.data_ptr())
    assert e.value.code == -1 and "device memory" in str(e.value)
    with pytest.raises(libipkgpu.IpkGpuError) as e:
        libipkgpu.best_prefix_noalloc(0, 0, a.data_ptr() + 4, 4, 8, 4, out.data_ptr())
    assert e.value.code == -1 and "aligned" in str(e.value)
