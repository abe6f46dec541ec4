// h2d_ceiling.cu -- the box's concurrent pinned host-to-device ceiling: one host thread per GPU, each
// with its own pinned buffer and stream, plain cudaMemcpyAsync of one large copy per iteration, all GPUs started together.
// Prints GB/s per GPU and aggregate for N = 1, 2, 4, 8 (as many as the box has), with the pinned buffer allocated (a) by
// the calling thread as is and (b) after binding the thread to the CPUs of the GPU's NUMA node (sysfs numa_node of the
// PCI device), so the effect of host-memory placement is visible.
// Build: nvcc -O2 -o h2d_ceiling h2d_ceiling.cu -lpthread
#ifndef _GNU_SOURCE
#define _GNU_SOURCE
#endif
#include <cuda_runtime.h>
#include <pthread.h>
#include <sched.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

static const size_t kBytes = 1ull << 30; // 1 GiB per copy
static const int kIters = 6;
static pthread_barrier_t bar;
struct Arg { int dev; int bind; double gbs; int node; };

static int numa_node_of(int dev)
{
    char bus[32], path[128];
    if (cudaDeviceGetPCIBusId(bus, sizeof bus, dev) != cudaSuccess) return -1;
    for (char *p = bus; *p; ++p) if (*p >= 'A' && *p <= 'Z') *p += 32;
    snprintf(path, sizeof path, "/sys/bus/pci/devices/%s/numa_node", bus);
    FILE *f = fopen(path, "r");
    int node = -1;
    if (f) { if (fscanf(f, "%d", &node) != 1) node = -1; fclose(f); }
    return node;
}
static void bind_to_node(int node)
{
    if (node < 0) return;
    char path[128], line[4096];
    snprintf(path, sizeof path, "/sys/devices/system/node/node%d/cpulist", node);
    FILE *f = fopen(path, "r");
    if (!f || !fgets(line, sizeof line, f)) { if (f) fclose(f); return; }
    fclose(f);
    cpu_set_t set; CPU_ZERO(&set);
    for (char *tok = strtok(line, ",\n"); tok; tok = strtok(NULL, ",\n")) {
        int a, b;
        if (sscanf(tok, "%d-%d", &a, &b) == 2) { for (int c = a; c <= b; ++c) CPU_SET(c, &set); }
        else if (sscanf(tok, "%d", &a) == 1) CPU_SET(a, &set);
    }
    sched_setaffinity(0, sizeof set, &set);
}
static void *worker(void *p)
{
    Arg *a = (Arg *)p;
    a->node = numa_node_of(a->dev);
    if (a->bind) bind_to_node(a->node);
    cudaSetDevice(a->dev);
    void *h = NULL, *d = NULL;
    cudaStream_t s;
    cudaEvent_t e0, e1;
    cudaMallocHost(&h, kBytes); memset(h, 1, kBytes); // first touch after binding
    cudaMalloc(&d, kBytes);
    cudaStreamCreate(&s); cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaMemcpyAsync(d, h, kBytes, cudaMemcpyHostToDevice, s); cudaStreamSynchronize(s);
    pthread_barrier_wait(&bar);
    cudaEventRecord(e0, s);
    for (int i = 0; i < kIters; ++i) cudaMemcpyAsync(d, h, kBytes, cudaMemcpyHostToDevice, s);
    cudaEventRecord(e1, s);
    cudaStreamSynchronize(s);
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    a->gbs = (double)kBytes * kIters / (ms * 1e-3) / 1e9;
    pthread_barrier_wait(&bar);
    cudaFreeHost(h); cudaFree(d);
    return NULL;
}
int main()
{
    int count = 0; cudaGetDeviceCount(&count);
    printf("{\"gpus\": %d, \"bytes_per_copy\": %zu, \"copies\": %d, \"runs\": [", count, kBytes, kIters);
    int first = 1;
    for (int bind = 0; bind < 2; ++bind)
        for (int n = 1; n <= count; n *= 2) {
            pthread_t th[16]; Arg arg[16];
            pthread_barrier_init(&bar, NULL, n);
            for (int i = 0; i < n; ++i) { arg[i].dev = i; arg[i].bind = bind; arg[i].gbs = 0; pthread_create(&th[i], NULL, worker, &arg[i]); }
            double sum = 0;
            for (int i = 0; i < n; ++i) { pthread_join(th[i], NULL); sum += arg[i].gbs; }
            pthread_barrier_destroy(&bar);
            printf("%s\n {\"n_gpus\": %d, \"numa_bound\": %d, \"aggregate_gbs\": %.1f, \"per_gpu_gbs\": [", first ? "" : ",", n, bind, sum);
            for (int i = 0; i < n; ++i) printf("%s%.1f", i ? ", " : "", arg[i].gbs);
            printf("], \"numa_nodes\": [");
            for (int i = 0; i < n; ++i) printf("%s%d", i ? ", " : "", arg[i].node);
            printf("]}");
            first = 0;
        }
    printf("\n]}\n");
    return 0;
}
