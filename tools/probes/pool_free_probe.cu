// Probe: which way of freeing a cudaMallocAsync block gives it back to the pool (cudaMemPoolAttrUsedMemCurrent)?
//   nvcc -arch=sm_100a -o /tmp/pool_free_probe tools/probes/pool_free_probe.cu && /tmp/pool_free_probe
#include <cstdio>
#include <cuda_runtime.h>
static unsigned long long used(cudaMemPool_t pool) { unsigned long long u = 0; cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &u); return u >> 20; }
int main()
{
  cudaSetDevice(0);
  cudaMemPool_t pool;
  cudaDeviceGetDefaultMemPool(&pool, 0);
  unsigned long long keep = ~0ull;
  cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
  for (int mode = 0; mode < 4; ++mode) {
    cudaStream_t s;
    cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
    void* p = nullptr;
    cudaMallocAsync(&p, 64u << 20, s);
    cudaMemsetAsync(p, 1, 64u << 20, s);
    cudaStreamSynchronize(s);
    const unsigned long long u0 = used(pool);
    cudaError_t rc = cudaSuccess;
    if (mode == 0) { rc = cudaFree(p); cudaStreamDestroy(s); }                       // cudaFree, stream alive
    if (mode == 1) { cudaStreamDestroy(s); rc = cudaFree(p); }                       // cudaFree after the stream is gone
    if (mode == 2) { cudaStreamDestroy(s); rc = cudaFreeAsync(p, 0); cudaStreamSynchronize(0); }   // legacy stream
    if (mode == 3) { rc = cudaFreeAsync(p, s); cudaStreamSynchronize(s); cudaStreamDestroy(s); }   // ordered on its stream
    std::printf("mode %d: rc=%d (%s) used %llu -> %llu MiB\n", mode, (int)rc, cudaGetErrorString(rc), u0, used(pool));
  }
  return 0;
}
