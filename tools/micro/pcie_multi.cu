// Probe for the end-to-end ceiling at N GPUs: every GPU copies its share of a 199 MB f64 frame to the host
// at the same time — into (a) its own cudaMallocHost buffer, (b) its rows of ONE registered malloc'd frame first touched by
// one thread (what a shared Canvas frame is), (c) the same frame with its pages interleaved over the NUMA nodes (mbind
// MPOL_INTERLEAVE before the first touch).  Prints the aggregate GB/s of each and the machine's topology hints.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pcie_multi pcie_multi.cu -lpthread ; run on a multi-GPU box.
#include <cuda_runtime.h>
#include <sys/mman.h>
#include <sys/syscall.h>
#include <unistd.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include <atomic>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main(int argc, char** argv) {
    int n = 0;
    CK(cudaGetDeviceCount(&n));
    if (argc > 1) n = std::min(n, atoi(argv[1]));
    const size_t frame = 3840ull * 2160 * 3 * 8;
    printf("GPUs: %d, frame %.1f MB\n", n, frame / 1e6);
    std::vector<void*> dev(n);
    std::vector<cudaStream_t> st(n);
    for (int g = 0; g < n; ++g) { CK(cudaSetDevice(g)); CK(cudaMalloc(&dev[g], frame)); CK(cudaMemset(dev[g], 1, frame)); CK(cudaStreamCreate(&st[g])); }
    for (int g = 0; g < n; ++g) { CK(cudaSetDevice(g)); CK(cudaDeviceSynchronize()); }
    auto run = [&](const char* name, std::vector<char*>& dst, size_t bytes_each, size_t pitch_dst, size_t tile, size_t tiles) {
        double best = 1e30;
        for (int rep = 0; rep < 4; ++rep) {
            std::atomic<int> ready{0};
            std::vector<std::thread> th;
            std::vector<double> t1(n);
            double t0 = 0;
            std::atomic<bool> go{false};
            for (int g = 0; g < n; ++g)
                th.emplace_back([&, g] {
                    cudaSetDevice(g);
                    ready.fetch_add(1);
                    while (!go.load()) {}
                    if (tiles) cudaMemcpy2DAsync(dst[g], pitch_dst, dev[g], tile, tile, tiles, cudaMemcpyDeviceToHost, st[g]);
                    else cudaMemcpyAsync(dst[g], dev[g], bytes_each, cudaMemcpyDeviceToHost, st[g]);
                    cudaStreamSynchronize(st[g]);
                    t1[g] = now();
                });
            while (ready.load() < n) {}
            t0 = now();
            go.store(true);
            for (auto& t : th) t.join();
            double end = 0;
            for (double x : t1) end = std::max(end, x);
            if (rep) best = std::min(best, end - t0);
        }
        printf("%-72s %7.3f ms  aggregate %6.1f GB/s  per GPU %5.1f GB/s\n", name, best * 1e3, bytes_each * n / best / 1e9, bytes_each / best / 1e9);
    };
    const size_t each = frame / n;
    {   // (a) own pinned buffers
        std::vector<char*> dst(n);
        for (int g = 0; g < n; ++g) { CK(cudaSetDevice(g)); CK(cudaMallocHost((void**)&dst[g], each)); memset(dst[g], 0, each); }
        run("(a) every GPU -> its own cudaMallocHost buffer, 1/N of the frame each", dst, each, 0, 0, 0);
        for (int g = 0; g < n; ++g) cudaFreeHost(dst[g]);
    }
    for (int mode = 0; mode < 2; ++mode) {   // (b) one frame, first touch by this thread; (c) interleaved over the NUMA nodes
        char* host = (char*)mmap(nullptr, frame, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
        if (host == MAP_FAILED) { printf("mmap failed\n"); return 1; }
        long rc = 0;
        if (mode == 1) {
            unsigned long mask = ~0ul;   // all nodes the kernel knows of (it intersects with the allowed set)
            rc = syscall(SYS_mbind, host, frame, 3 /* MPOL_INTERLEAVE */, &mask, 64ul, 0u);
            if (rc != 0) { unsigned long m2 = 3ul; rc = syscall(SYS_mbind, host, frame, 3, &m2, 3ul, 0u); }
        }
        memset(host, 0, frame);
        CK(cudaHostRegister(host, frame, cudaHostRegisterPortable));
        std::vector<char*> dst(n);
        char name[160];
        for (int g = 0; g < n; ++g) dst[g] = host + (size_t)g * each;
        snprintf(name, sizeof name, "(%c) one registered frame%s: contiguous 1/N each", mode ? 'c' : 'b', mode ? (rc == 0 ? ", pages interleaved over NUMA nodes" : ", mbind FAILED") : ", first touch by one thread");
        run(name, dst, each, 0, 0, 0);
        const size_t tile = 4 * 3840ull * 3 * 8, tiles = 2160 / 4 / n;
        for (int g = 0; g < n; ++g) dst[g] = host + (size_t)g * tile;
        snprintf(name, sizeof name, "(%c) same frame: interleaved 4-row tiles (pitched 2-D DMA, %zu tiles each)", mode ? 'c' : 'b', tiles);
        run(name, dst, tile * tiles, tile * n, tile, tiles);
        cudaHostUnregister(host);
        munmap(host, frame);
    }
    return 0;
}
