/* Plain C caller of the data plane: one rank (loop-back), four rounds through the native round engine.
 *
 *   gcc -Iinclude examples/native_round.c -o build/native_round akka_allreduce_b200/libaar_cuda.so -lcudart -Wl,-rpath,$PWD/akka_allreduce_b200
 *
 * With more ranks the regions[] table holds every peer's region as mapped into this process (aar_vmm_* for one process per
 * GPU, see akka_allreduce_b200/parallel/symmetric.py for the exchange).  Exits 0 on success, 3 when no GPU is usable. */
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#include "akka_allreduce_b200.h"

#define CHECK(call)                                                       \
  do {                                                                    \
    int rc_ = (call);                                                     \
    if (rc_ != 0) {                                                       \
      fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, aar_last_error()); \
      return 3;                                                           \
    }                                                                     \
  } while (0)

int main(void) {
  const int n = 1 << 20, chunk = 1 << 16, rounds = 4;
  AarLayout lay;
  CHECK(aar_layout_compute(1, n, chunk, /*maxLag=*/1, AAR_F32, /*oneshot=*/0, /*stage=*/0, &lay));
  void* region = NULL;
  CHECK(aar_region_alloc_local(lay.regionBytes, &region));
  void* regions[1] = {region};
  AarEngine* eng = NULL;
  CHECK(aar_engine_create(&lay, 0, regions, NULL, /*minRequired=*/1, lay.totalChunks, 0, 0ull, /*algo=*/0, /*grid=*/0, /*overlap=*/0, &eng));

  float *in = NULL, *out = NULL, *host = (float*)malloc(sizeof(float) * n);
  int* counts = NULL;
  if (cudaMalloc((void**)&in, sizeof(float) * n) || cudaMalloc((void**)&out, sizeof(float) * n) || cudaMalloc((void**)&counts, sizeof(int) * lay.totalChunks)) return 3;
  int bad = 0;
  for (int r = 0; r < rounds; ++r) {
    for (int i = 0; i < n; ++i) host[i] = (float)(i + r); /* the reference's demo source: data[i] = i + iteration */
    cudaMemcpy(in, host, sizeof(float) * n, cudaMemcpyHostToDevice);
    CHECK(aar_engine_enqueue(eng, r, in, out, counts, 0, /*stream=*/NULL));
    CHECK(aar_engine_synchronize(eng));
    cudaMemcpy(host, out, sizeof(float) * n, cudaMemcpyDeviceToHost);
    for (int i = 0; i < n; ++i) bad += host[i] != (float)(i + r);
  }
  printf("native_round: %d rounds flushed, %d mismatches, %lld launches\n", aar_engine_rounds_flushed(eng), bad, aar_engine_launches(eng));
  CHECK(aar_engine_destroy(eng));
  CHECK(aar_region_free_local(region));
  cudaFree(in);
  cudaFree(out);
  cudaFree(counts);
  free(host);
  return bad ? 1 : 0;
}
