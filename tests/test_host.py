"""Host side above the device ABI (ipk_b200/host): the raxml-ng posterior loader against the REFERENCE's own csv
reader (oracle/_ref/ref_csv_*, built from /root/reference's csv.h as ar.cpp instantiates it) and against committed
fixtures of it; the .ipk writer/reader round trip; and -- on a GPU -- the ipk-b200 executable end to end."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
REF_CSV = {4: os.path.join(ROOT, "oracle", "_ref", "ref_csv_dna"), 20: os.path.join(ROOT, "oracle", "_ref", "ref_csv_aa")}
AA_RAXML = "ARNDCQEGHILKMFPSTWYV"


def write_tsv(path, probs, labels, sigma, rng=None, tricky=False):
    """probs [n, W, sigma] in the FILE's column order; numbers printed the way raxml-ng does (--precision 9), plus a few
    other spellings the csv parser accepts when `tricky`"""
    names = "ACGT" if sigma == 4 else AA_RAXML
    with open(path, "w") as f:
        f.write("Node\tSite\tState\t" + "\t".join("p_" + c for c in names) + "\n")
        for n, lab in enumerate(labels):
            for s in range(probs.shape[1]):
                row = probs[n, s]
                cells = []
                for j, p in enumerate(row):
                    t = f"{p:.9f}"
                    if tricky and rng is not None:
                        r = rng.integers(0, 12)
                        if r == 0: t = f"{p:.6e}"
                        elif r == 1: t = f" {p:.5f} "
                        elif r == 2: t = f"{p:.17f}"
                        elif r == 3 and p < 1e-3: t = f"{p:.3E}"
                        elif r == 4: t = "+" + t
                    cells.append(t)
                f.write(f"{lab}\t{s + 1}\t{names[int(np.argmax(row))]}\t" + "\t".join(cells) + "\n")
            if tricky and n == 1:
                f.write("\n.a comment line as single_and_empty_line_comment<'.'> skips it\n")


def random_probs(rng, n, W, sigma):
    p = rng.dirichlet(np.full(sigma, 0.3), size=(n, W))
    p[rng.random((n, W)) < 0.1] = np.eye(sigma)[0]                       # exact 1 / 0 columns -> log10(0) = -inf
    return p


def ref_parse(tsv, sigma, tmp_path):
    out = str(tmp_path / "ref.f32")
    subprocess.run([REF_CSV[sigma], tsv, out], check=True)
    return np.fromfile(out, np.float32)


@pytest.mark.parametrize("sigma", [4, 20])
def test_loader_matches_reference_csv_reader(sigma, tmp_path):
    if not os.path.exists(REF_CSV[sigma]):
        pytest.skip("oracle/_ref/ref_csv_* not built (needs /root/reference)")
    from ipk_b200 import hostlib
    rng = np.random.default_rng(3 + sigma)
    labels = [f"Node{i}" for i in range(1, 8)]
    probs = random_probs(rng, len(labels), 37, sigma)
    tsv = str(tmp_path / "probs.tsv")
    write_tsv(tsv, probs, labels, sigma, rng, tricky=True)
    want = ref_parse(tsv, sigma, tmp_path).reshape(len(labels), 37, sigma)
    got_labels, got = hostlib.read_ar(tsv, sigma)
    assert got_labels == labels
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), "loader differs bitwise from the reference's csv reader"
    # a subset, in another order, as db_builder asks for nodes
    pick = ["Node5", "Node2", "Node7"]
    _, sub = hostlib.read_ar(tsv, sigma, pick, threads=2)
    assert np.array_equal(sub.view(np.uint32), want[[4, 1, 6]].view(np.uint32))


@pytest.mark.parametrize("sigma", [4, 20])
def test_loader_matches_committed_fixture(sigma):
    """the same comparison against outputs of the reference's reader committed under tests/golden (make_golden_ar.py)"""
    from ipk_b200 import hostlib
    tsv = os.path.join(GOLD, f"ar_{'dna' if sigma == 4 else 'aa'}.tsv")
    want = np.fromfile(os.path.join(GOLD, f"ar_{'dna' if sigma == 4 else 'aa'}.f32"), np.float32)
    labels, got = hostlib.read_ar(tsv, sigma)
    assert len(labels) == 3
    assert np.array_equal(got.reshape(-1).view(np.uint32), want.view(np.uint32))


def test_parse_float_known_answers():
    """csv.h:1024-1085 accumulates in float: not the correctly rounded value for long digit strings"""
    from ipk_b200 import hostlib
    f = np.float32
    assert hostlib.parse_float("0.5") == 0.5 and hostlib.parse_float("1") == 1.0 and hostlib.parse_float("-2.25") == -2.25
    x = f(0)
    pos = f(1)
    for ch in "123456789":                    # 0.123456789 the reference's way
        pos = f(pos / f(10))
        x = f(x + f(int(ch) * pos))
    assert hostlib.parse_float("0.123456789") == float(x)
    # 1e-3: base 0.1f, e = 3 -> x *= base (e = 2); base = base*base (e = 1); x *= base   (csv.h:1058-1070)
    b = f(0.1)
    assert hostlib.parse_float("1e-3") == float(f(f(f(1) * b) * f(b * b)))
    with pytest.raises(hostlib.HostError):
        hostlib.parse_float("0.5x")
    with pytest.raises(hostlib.HostError):
        hostlib.parse_float("abc")


def test_loader_errors(tmp_path):
    from ipk_b200 import hostlib
    with pytest.raises(hostlib.HostError):
        hostlib.read_ar(str(tmp_path / "missing.tsv"), 4)
    tsv = str(tmp_path / "p.tsv")
    write_tsv(tsv, random_probs(np.random.default_rng(1), 2, 5, 4), ["A1", "A2"], 4)
    with pytest.raises(hostlib.HostError, match="Could not read the AR matrix"):
        hostlib.read_ar(tsv, 4, ["A1", "nope"])
    with open(tsv, "a") as f:
        f.write("A3\t1\tA\t0.5\t0.5\n")       # too few columns
    with pytest.raises(hostlib.HostError):
        hostlib.read_ar(tsv, 4, ["A3"])


def test_ipk_round_trip(tmp_path):
    from ipk_b200 import hostlib
    rng = np.random.default_rng(9)
    nk = 1000
    cnt = rng.integers(1, 9, nk)
    offsets = np.concatenate(([0], np.cumsum(cnt))).astype(np.uint64)
    ne = int(offsets[-1])
    keys = rng.permutation(1 << 16)[:nk].astype(np.uint64)
    fv = np.sort(rng.random(nk).astype(np.float32))
    branch = rng.integers(0, 500, ne).astype(np.uint32)
    score = (-5 * rng.random(ne)).astype(np.float32)
    path = str(tmp_path / "DB.ipk")
    hostlib.write_ipk(path, keys, fv, offsets, branch, score, 8, 1.5, "DNA", "((a:1,b:2):0.5,c:3);", [1, 1, 3, 1, 5], [1, 2, 3.5, 3, 6.5])
    d = hostlib.read_ipk(path)
    assert d["sequence_type"] == "DNA" and d["tree"] == "((a:1,b:2):0.5,c:3);" and d["kmer_size"] == 8 and d["omega"] == 1.5
    assert list(d["index_num_nodes"]) == [1, 1, 3, 1, 5] and list(d["index_len"]) == [1, 2, 3.5, 3, 6.5]
    for name, a in (("keys", keys), ("fv", fv), ("offsets", offsets), ("branch", branch), ("score", score)):
        assert np.array_equal(d[name], a), name
    assert os.path.getsize(path) == 8 + 22 + 2 + (8 + 3) + 8 + 5 * 16 + (8 + 20) + 8 + 4 + 8 + 8 + nk * 20 + ne * 8
    with open(path, "r+b") as f:              # truncated file -> format error, not a crash
        f.truncate(os.path.getsize(path) - 5)
    with pytest.raises(hostlib.HostError):
        hostlib.read_ipk(path)
    hostlib.write_ipk(path, [], [], [0], [], [], 8, 1.5)
    assert len(hostlib.read_ipk(path)["keys"]) == 0


@pytest.mark.gpu
@pytest.mark.parametrize("sigma,bits,k,omega,filt", [(4, 2, 8, 1.5, "--mif0"), (20, 5, 3, 1.5, "--random")])
def test_ipk_b200_executable_end_to_end(sigma, bits, k, omega, filt, tmp_path):
    """AR file + node map -> ipk-b200 -> DB.ipk, compared entry for entry with the oracle build on the loader's matrices"""
    import oraclelib as O
    from parity import assert_db_equal
    from types import SimpleNamespace
    from ipk_b200 import hostlib, capi, synth
    n_taxa, W = 6, 60
    logp, nb = synth.posteriors(n_taxa, W, sigma, seed=5)
    probs = np.power(10.0, logp.astype(np.float64))
    if sigma == 20:                            # the file is in raxml-ng's column order; logp is in IPK's
        order = "RHKDESTNQCGPAILMFWYV"
        probs = probs[:, :, [order.index(c) for c in AA_RAXML]]
    labels = [f"Node{i + 1}" for i in range(len(nb))]
    tsv, nmap, out = str(tmp_path / "ar.raxml.ancestralProbs"), str(tmp_path / "node_map.tsv"), str(tmp_path / "DB.ipk")
    write_tsv(tsv, probs, labels, sigma)
    with open(nmap, "w") as f:
        f.write("# AR node label <TAB> original branch (post-order id)\n")
        for lab, b in zip(labels, nb):
            f.write(f"{lab}\t{int(b)}\n")
    r = subprocess.run([hostlib.EXE_PATH, "--ar-probs", tsv, "--node-map", nmap, "--states", "dna" if sigma == 4 else "aa", "-k", str(k),
                        "--omega", str(omega), filt, "-o", out, "--node-count", str(2 * n_taxa - 1)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    d = hostlib.read_ipk(out)
    got = SimpleNamespace(keys=d["keys"], fv=d["fv"], offsets=d["offsets"], branch=d["branch"], score=d["score"],
                          n_kmers=len(d["keys"]), n_entries=len(d["branch"]))
    _, parsed = hostlib.read_ar(tsv, sigma, labels)
    thr, lthr = capi.score_threshold(omega, sigma, k)
    want = O.build(parsed, nb, W, sigma, bits, k, lthr, thr, 2 * n_taxa - 1, 0 if filt == "--mif0" else 1, 2)
    assert want.n_entries > 100
    assert_db_equal(got, want)
    assert d["kmer_size"] == k and d["sequence_type"] == ("DNA" if sigma == 4 else "Proteins")
    # error convention of the reference binary: message on stderr, exit status 1 (main.cpp:225-246)
    r = subprocess.run([hostlib.EXE_PATH, "--ar-probs", tsv + ".missing", "--node-map", nmap], capture_output=True, text=True)
    assert r.returncode == 1 and "Exception" in r.stderr
