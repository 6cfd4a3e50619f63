// nsgpu.cu -- single translation unit of libnsgpu.so (B200 / sm_100a).
// Build: nvcc -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -shared nsgpu.cu -lnccl
#include <algorithm>
#include <cstdint>
#include <mutex>
#include "ns_ctx.cuh"
#include "ns_setup.inl"
#include "ns_solver.inl"
#include "ns_api.inl"
#include "ns_noalloc.inl"
