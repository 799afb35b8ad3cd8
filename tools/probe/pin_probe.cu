// pin_probe.cu -- how fast can 8 GB of page-locked host memory be had on a cold process?  (tools/probe; not product code)
//   cudaMallocHost  vs  anonymous mmap + MADV_HUGEPAGE + parallel first touch + cudaHostRegister
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>
#include <sys/mman.h>
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main(int argc, char** argv)
{
    const size_t bytes = size_t(argc > 1 ? atof(argv[1]) : 8.0) * (1ull << 30);
    cudaFree(0);
    double t = now();
    void* a = nullptr;
    cudaError_t e = cudaMallocHost(&a, bytes);
    printf("cudaMallocHost %.1f GiB: %.3f s (%s)\n", bytes / double(1ull << 30), now() - t, cudaGetErrorString(e));
    cudaFreeHost(a);
    for (int huge = 0; huge < 2; ++huge)
        for (int threads : { 1, 16 })
        {
            t = now();
            void* p = mmap(nullptr, bytes + (2 << 20), PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
            char* q = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(p) + (2 << 20) - 1) & ~uintptr_t((2 << 20) - 1));
            if (huge) madvise(q, bytes, MADV_HUGEPAGE);
            std::vector<std::thread> th;
            for (int i = 0; i < threads; ++i)
                th.emplace_back([=] { const size_t lo = bytes / threads * i, hi = i + 1 == threads ? bytes : bytes / threads * (i + 1); for (size_t o = lo; o < hi; o += 4096) q[o] = 0; });
            for (auto& x : th) x.join();
            const double t1 = now();
            e = cudaHostRegister(q, bytes, cudaHostRegisterDefault);
            printf("mmap%s + touch by %2d threads: %.3f s, cudaHostRegister: %.3f s (%s)\n", huge ? " + MADV_HUGEPAGE" : "", threads, t1 - t, now() - t1, cudaGetErrorString(e));
            if (e == cudaSuccess) cudaHostUnregister(q);
            munmap(p, bytes + (2 << 20));
        }
    return 0;
}
