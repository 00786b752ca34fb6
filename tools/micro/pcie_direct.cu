// Probe for the direct Canvas delivery of rtc_render: how fast does a finished frame reach a caller
// buffer that was registered once (cudaHostRegister) — by DMA (1-D and pitched 2-D copies of interleaved row tiles)
// and by a widening kernel that stores f64 straight into the mapped buffer — against the staged f32 copy.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pcie_direct pcie_direct.cu -lrt ; run on a GPU box.
#include <cuda_runtime.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <unistd.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__global__ void widen_store(const float4* __restrict__ src, double2* __restrict__ dst, size_t n4) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 v = src[i];
        dst[2 * i] = make_double2((double)v.x, (double)v.y);
        dst[2 * i + 1] = make_double2((double)v.z, (double)v.w);
    }
}
__global__ void copy_store(const double2* __restrict__ src, double2* __restrict__ dst, size_t n2) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}

int main() {
    const size_t W = 3840, H = 2160, n = W * H * 3, b64 = n * 8, b32 = n * 4;
    double* d64; float* d32;
    CK(cudaMalloc(&d64, b64)); CK(cudaMalloc(&d32, b32));
    CK(cudaMemset(d64, 0, b64)); CK(cudaMemset(d32, 0, b32));
    cudaStream_t st; CK(cudaStreamCreate(&st));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    auto timed = [&](const char* name, size_t bytes, auto fn) {
        float best = 1e30f;
        for (int rep = 0; rep < 5; ++rep) {
            cudaEventRecord(e0, st); fn(); cudaEventRecord(e1, st); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep) best = best < ms ? best : ms;
        }
        cudaError_t e = cudaGetLastError();
        printf("%-64s %8.3f ms  %7.1f GB/s %s\n", name, best, bytes / (best * 1e-3) / 1e9, e == cudaSuccess ? "" : cudaGetErrorString(e));
    };
    // (a) pinned staging (cudaMallocHost), f32: today's path
    float* stage; CK(cudaMallocHost(&stage, b32));
    timed("f32 frame -> cudaMallocHost staging (1-D DMA)", b32, [&] { cudaMemcpyAsync(stage, d32, b32, cudaMemcpyDeviceToHost, st); });
    // (b) caller memory: malloc + cudaHostRegister
    for (int kind = 0; kind < 2; ++kind) {
        void* host = nullptr; int fd = -1;
        if (kind == 0) { if (posix_memalign(&host, 4096, b64)) return 1; }
        else {
            fd = shm_open("/rtc_probe_shm", O_CREAT | O_RDWR, 0600);
            if (fd < 0 || ftruncate(fd, (off_t)b64)) { printf("shm failed\n"); return 1; }
            host = mmap(nullptr, b64, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
            if (host == MAP_FAILED) { printf("mmap failed\n"); return 1; }
        }
        const char* kn = kind == 0 ? "malloc" : "POSIX shm";
        auto t0 = std::chrono::steady_clock::now();
        memset(host, 0, b64);
        auto t1 = std::chrono::steady_clock::now();
        cudaError_t er = cudaHostRegister(host, b64, cudaHostRegisterPortable | cudaHostRegisterMapped);
        auto t2 = std::chrono::steady_clock::now();
        printf("[%s] first touch %.1f ms, cudaHostRegister(199 MB) %.1f ms: %s\n", kn, std::chrono::duration<double, std::milli>(t1 - t0).count(),
               std::chrono::duration<double, std::milli>(t2 - t1).count(), cudaGetErrorString(er));
        if (er != cudaSuccess) continue;
        void* mapped = nullptr; CK(cudaHostGetDevicePointer(&mapped, host, 0));
        char name[128];
        snprintf(name, sizeof name, "[%s] f64 frame -> registered buffer (1-D DMA)", kn);
        timed(name, b64, [&] { cudaMemcpyAsync(host, d64, b64, cudaMemcpyDeviceToHost, st); });
        snprintf(name, sizeof name, "[%s] f32 frame -> registered buffer (1-D DMA)", kn);
        timed(name, b32, [&] { cudaMemcpyAsync(host, d32, b32, cudaMemcpyDeviceToHost, st); });
        for (int shards : {2, 4, 8}) {   // one shard's interleaved 4-row tiles: compact device rows -> frame rows
            const size_t tile = 4 * W * 3 * 8, tiles = H / 4 / shards;
            snprintf(name, sizeof name, "[%s] f64 shard 1/%d, 4-row tiles (2-D DMA, %zu x %zu KB)", kn, shards, tiles, tile >> 10);
            timed(name, tile * tiles, [&] { cudaMemcpy2DAsync(host, tile * shards, d64, tile, tile, tiles, cudaMemcpyDeviceToHost, st); });
        }
        for (int bands : {8}) {
            snprintf(name, sizeof name, "[%s] f64 frame in %d bands (1-D DMA each)", kn, bands);
            timed(name, b64, [&] { for (int b = 0; b < bands; ++b) cudaMemcpyAsync((char*)host + b * (b64 / bands), (char*)d64 + b * (b64 / bands), b64 / bands, cudaMemcpyDeviceToHost, st); });
        }
        snprintf(name, sizeof name, "[%s] widen kernel f32 HBM -> f64 mapped stores (16 B/thread)", kn);
        timed(name, b64, [&] { widen_store<<<148 * 8, 256, 0, st>>>((const float4*)d32, (double2*)mapped, n / 4); });
        snprintf(name, sizeof name, "[%s] copy kernel f64 HBM -> f64 mapped stores (16 B/thread)", kn);
        timed(name, b64, [&] { copy_store<<<148 * 8, 256, 0, st>>>((const double2*)d64, (double2*)mapped, n / 2); });
        snprintf(name, sizeof name, "[%s] widen kernel, 1/8 of the frame", kn);
        timed(name, b64 / 8, [&] { widen_store<<<148 * 8, 256, 0, st>>>((const float4*)d32, (double2*)mapped, n / 32); });
        cudaHostUnregister(host);
        if (kind == 1) { munmap(host, b64); close(fd); shm_unlink("/rtc_probe_shm"); } else free(host);
    }
    return 0;
}
