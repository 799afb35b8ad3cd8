"""ipk_b200 -- B200-native (sm_100a) phylo-k-mer database build: one hot path of phylo42/IPK (ipk::build,
ipk/include/db_builder.h:22-29) behind a C ABI (include/ipkgpu.h).

  ipk_b200/csrc/      CUDA kernels + the C ABI (libipkgpu.so), synthetic-input generator (libipksynth.so)
  ipk_b200/host/      C++ host mirroring the reference's ipk-dna / ipk-aa command line and ipk::build signature
  ipk_b200/capi       ctypes binding of the C ABI (harness plumbing for tests / bench / multi-GPU runs)
"""
__all__ = ["capi", "synth"]
