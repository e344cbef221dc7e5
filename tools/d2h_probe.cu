// d2h_probe -- what can this box's host side sustain when N GPUs copy their point clouds out at once?
//
// VERDICT r01 (weak #7): the end-to-end rate stops scaling beyond one GPU (per-GPU D2H falls from 51.7 GB/s to 11.2 GB/s
// at 8 GPUs) and nobody had measured the box's ceiling.  This probe copies `bytes` (default 469 MB = one frame of 16-byte
// XYZI clouds of the bench, in 64 pieces like the bench does) device -> pinned host on N GPUs concurrently, one host
// thread per GPU, for N = 1, 2, 4, 8 (up to the visible devices), with the pinned buffers
//   plain : cudaHostAlloc'ed by the main thread (wherever the kernel puts them -- what torch's pin_memory() gives)
//   numa  : lsb200_alloc_host (thread bound to the CPUs of the GPU's NUMA node for allocation + first touch)
// and prints one JSON line per (mode, N): per-GPU and aggregate GB/s (max-over-GPUs time).
//
// build: make -C lidarshooter_b200 probe        run: tools/d2h_probe.bin [bytes] [pieces] [reps]
#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

#include "../include/lsb200.h"

static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main(int argc, char **argv)
{
    const size_t bytes = argc > 1 ? (size_t)atoll(argv[1]) : (size_t)469155488;
    const int pieces = argc > 2 ? atoi(argv[2]) : 64;
    const int reps = argc > 3 ? atoi(argv[3]) : 8;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) { fprintf(stderr, "no CUDA device\n"); return 1; }
    std::vector<lsb200_ctx *> ctx(n_dev, nullptr);
    std::vector<void *> d_buf(n_dev, nullptr), h_plain(n_dev, nullptr), h_numa(n_dev, nullptr);
    std::vector<cudaStream_t> st(n_dev);
    const size_t piece = (bytes / pieces) & ~(size_t)255;
    for (int g = 0; g < n_dev; ++g) {
        if (lsb200_create(g, &ctx[g]) != LSB200_OK) { fprintf(stderr, "lsb200_create(%d) failed\n", g); return 1; }
        cudaSetDevice(g);
        cudaMalloc(&d_buf[g], bytes);
        cudaMemset(d_buf[g], 1, bytes);
        cudaStreamCreateWithFlags(&st[g], cudaStreamNonBlocking);
        cudaHostAlloc(&h_plain[g], bytes, cudaHostAllocPortable);                  // main thread, no placement
        if (lsb200_alloc_host(ctx[g], bytes, &h_numa[g]) != LSB200_OK) { fprintf(stderr, "alloc_host failed\n"); return 1; }
        printf("{\"gpu\": %d, \"numa_node\": %d}\n", g, lsb200_device_numa_node(ctx[g]));
    }
    cudaDeviceSynchronize();
    for (int mode = 0; mode < 2; ++mode) {
        for (int n = 1; n <= n_dev; n *= 2) {
            std::vector<double> secs(n, 0.0);
            std::atomic<int> ready{0};
            std::atomic<bool> go{false};
            std::vector<std::thread> th;
            for (int g = 0; g < n; ++g)
                th.emplace_back([&, g]() {
                    cudaSetDevice(g);
                    char *h = static_cast<char *>(mode ? h_numa[g] : h_plain[g]);
                    char *d = static_cast<char *>(d_buf[g]);
                    for (int p = 0; p < pieces; ++p) cudaMemcpyAsync(h + p * piece, d + p * piece, piece, cudaMemcpyDeviceToHost, st[g]);   // warm-up
                    cudaStreamSynchronize(st[g]);
                    ready.fetch_add(1);
                    while (!go.load()) std::this_thread::yield();
                    const double t0 = now_s();
                    for (int r = 0; r < reps; ++r)
                        for (int p = 0; p < pieces; ++p) cudaMemcpyAsync(h + p * piece, d + p * piece, piece, cudaMemcpyDeviceToHost, st[g]);
                    cudaStreamSynchronize(st[g]);
                    secs[g] = now_s() - t0;
                });
            while (ready.load() < n) std::this_thread::yield();
            go.store(true);
            for (auto &t : th) t.join();
            double worst = 0.0;
            for (double s : secs) worst = s > worst ? s : worst;
            const double per_copy = (double)piece * pieces * reps;
            printf("{\"probe\": \"d2h\", \"pinned\": \"%s\", \"gpus\": %d, \"bytes_per_gpu_per_rep\": %zu, \"pieces\": %d, \"reps\": %d, "
                   "\"aggregate_gbs\": %.2f, \"per_gpu_gbs\": [", mode ? "numa (lsb200_alloc_host)" : "plain (cudaHostAlloc, main thread)", n,
                   piece * pieces, pieces, reps, n * per_copy / worst / 1e9);
            for (int g = 0; g < n; ++g) printf("%s%.2f", g ? ", " : "", per_copy / secs[g] / 1e9);
            printf("]}\n");
            fflush(stdout);
        }
    }
    for (int g = 0; g < n_dev; ++g) {
        cudaSetDevice(g);
        cudaFree(d_buf[g]); cudaFreeHost(h_plain[g]); lsb200_free_host(ctx[g], h_numa[g]); lsb200_destroy(ctx[g]);
    }
    return 0;
}
