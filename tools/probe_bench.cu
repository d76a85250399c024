// Standalone timing harness for the q5-9 match-finding kernels (development tool, not part of the library):
//   nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -o /tmp/probe_bench tools/probe_bench.cu
//   /tmp/probe_bench input.bin [depth] [bucket_bits]
// Cuts the file into 4 MiB shards, runs the two sort passes of the library, then every probe variant 5 times
// (CUDA events), and compares each variant's best[] with k_probe_sorted's.
#include <cstdio>
#include <algorithm>
#include "../brotli_lib_b200/csrc/kernels.cuh"
#include "probe_variants.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
using namespace bgpu;

#define CKX(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void k_checksum(const uint32_t* a, size_t n, unsigned long long* out) {
  unsigned long long s = 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) s += (unsigned long long)a[i] * (2 * i + 1);
  atomicAdd(out, s);
}

int main(int argc, char** argv) {
  if (argc < 2) { printf("usage: probe_bench file [depth] [bucket_bits]\n"); return 1; }
  const int depth = argc > 2 ? atoi(argv[2]) : 16, bb = argc > 3 ? atoi(argv[3]) : 14;
  FILE* f = fopen(argv[1], "rb"); if (!f) { printf("cannot open %s\n", argv[1]); return 1; }
  fseek(f, 0, SEEK_END); size_t n = ftell(f); fseek(f, 0, SEEK_SET);
  std::vector<uint8_t> h(n + 1024, 0); if (fread(h.data(), 1, n, f) != n) return 1; fclose(f);
  const uint32_t SH = 1u << 22; const uint32_t n_shard = (uint32_t)((n + SH - 1) / SH);
  std::vector<ShardD> shards(n_shard); uint32_t stile = 0, max_len = 0;
  for (uint32_t k = 0; k < n_shard; k++) {
    ShardD s = {}; s.in_off = (uint64_t)k * SH; s.len = (uint32_t)std::min<size_t>(SH, n - s.in_off); s.first_stile = stile;
    stile += (s.len - 3 + SORT_TILE - 1) / SORT_TILE; max_len = std::max(max_len, s.len); shards[k] = s;
  }
  uint8_t* d_in; ShardD* d_sh; uint32_t *sa_pos, *sa_v, *sb_pos, *sb_v, *hist, *best, *ref; unsigned long long* d_sum;
  CKX(cudaMalloc(&d_in, n + 1024)); CKX(cudaMemcpy(d_in, h.data(), n + 1024, cudaMemcpyHostToDevice));
  CKX(cudaMalloc(&d_sh, n_shard * sizeof(ShardD))); CKX(cudaMemcpy(d_sh, shards.data(), n_shard * sizeof(ShardD), cudaMemcpyHostToDevice));
  for (uint32_t** p : {&sa_pos, &sa_v, &sb_pos, &sb_v, &best, &ref}) CKX(cudaMalloc(p, (n + 1024) * 4));
  CKX(cudaMalloc(&hist, ((size_t)stile + 1) * SORT_BINS * 4)); CKX(cudaMalloc(&d_sum, 8));
  const uint32_t max_npos = max_len - 3;
  const dim3 gs((max_npos + SORT_TILE - 1) / SORT_TILE, n_shard);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto timed = [&](const char* name, auto&& fn, int reps = 5) {
    fn(); CKX(cudaDeviceSynchronize());
    cudaEventRecord(e0);
    for (int i = 0; i < reps; i++) fn();
    cudaEventRecord(e1); CKX(cudaEventSynchronize(e1));
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("%-28s %8.3f ms  (%.2f ms per GiB)\n", name, ms / reps, ms / reps * (double)(1u << 30) / n);
  };
  const int nb0 = 1 << SORT_LOW_BITS, nb1 = 1 << (bb - SORT_LOW_BITS);
  timed("sort pass 0", [&] {
    k_sort_hist<0><<<gs, 256>>>(d_in, nullptr, d_sh, hist, bb, 0, SORT_LOW_BITS);
    k_sort_scan<<<n_shard, 1024>>>(d_sh, hist, nb0);
    k_sort_scatter<0><<<gs, 256, SORT_SCATTER_SMEM>>>(d_in, nullptr, nullptr, d_sh, hist, sa_pos, sa_v, bb, 0, SORT_LOW_BITS);
  });
  timed("sort pass 1", [&] {
    k_sort_hist<1><<<gs, 256>>>(d_in, sa_v, d_sh, hist, bb, SORT_LOW_BITS, bb - SORT_LOW_BITS);
    k_sort_scan<<<n_shard, 1024>>>(d_sh, hist, nb1);
    k_sort_scatter<1><<<gs, 256, SORT_SCATTER_SMEM>>>(d_in, sa_pos, sa_v, d_sh, hist, sb_pos, sb_v, bb, SORT_LOW_BITS, bb - SORT_LOW_BITS);
  });
  auto sort_variant = [&](const char* name, auto k0, auto k1, int threads) {
    cudaFuncSetAttribute(k0, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536); cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    uint32_t* tp = best; uint32_t* tv = ref;   // scratch outputs of pass 1 (compared with sb_pos / sb_v below)
    timed(name, [&] {
      k_sort_hist<0><<<gs, 256>>>(d_in, nullptr, d_sh, hist, bb, 0, SORT_LOW_BITS);
      k_sort_scan<<<n_shard, 1024>>>(d_sh, hist, nb0);
      k0<<<gs, threads, SORT_SCATTER_SMEM>>>(d_in, nullptr, nullptr, d_sh, hist, sa_pos, sa_v, bb, 0, SORT_LOW_BITS);
      k_sort_hist<1><<<gs, 256>>>(d_in, sa_v, d_sh, hist, bb, SORT_LOW_BITS, bb - SORT_LOW_BITS);
      k_sort_scan<<<n_shard, 1024>>>(d_sh, hist, nb1);
      k1<<<gs, threads, SORT_SCATTER_SMEM>>>(d_in, sa_pos, sa_v, d_sh, hist, tp, tv, bb, SORT_LOW_BITS, bb - SORT_LOW_BITS);
    });
    unsigned long long s[4]; const uint32_t* arr[4] = {sb_pos, tp, sb_v, tv};
    for (int k = 0; k < 4; k++) { cudaMemset(d_sum, 0, 8); k_checksum<<<1024, 256>>>(arr[k], n - 3, d_sum); cudaMemcpy(&s[k], d_sum, 8, cudaMemcpyDeviceToHost); }
    printf("    %s: %s\n", name, s[0] == s[1] && s[2] == s[3] ? "same order" : "DIFFERENT order");
  };
  sort_variant("sort both passes, 8 warps", k_sort_scatter<0, 8>, k_sort_scatter<1, 8>, 256);
  sort_variant("sort both passes, 16 warps", k_sort_scatter<0, 16>, k_sort_scatter<1, 16>, 512);
  sort_variant("sort both passes, 32 warps", k_sort_scatter<0, 32>, k_sort_scatter<1, 32>, 1024);
  auto check = [&](const char* name, const uint32_t* b) {
    unsigned long long s[2];
    for (int k = 0; k < 2; k++) {
      cudaMemset(d_sum, 0, 8); k_checksum<<<1024, 256>>>(k ? b : ref, n, d_sum); cudaMemcpy(&s[k], d_sum, 8, cudaMemcpyDeviceToHost);
    }
    printf("    %s: %s\n", name, s[0] == s[1] ? "same best[]" : "DIFFERENT best[]");
  };
  const dim3 gp((max_npos + PROBE_TILE - 1) / PROBE_TILE, n_shard);
  CKX(cudaFuncSetAttribute(k_probe_sorted, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PROBE_SMEM));
  timed("k_probe_sorted", [&] { k_probe_sorted<<<gp, 256, PROBE_SMEM>>>(d_in, d_sh, sb_pos, sb_v, ref, depth, bb); });
  const size_t sm = (size_t)(PROBE_TILE + depth) * 24;
  CKX(cudaFuncSetAttribute(k_probe_diag<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
  CKX(cudaFuncSetAttribute(k_probe_diag<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
  CKX(cudaMemset(best, 0xFF, n * 4));
  timed("k_probe_diag", [&] { k_probe_diag<false><<<gp, 256, sm>>>(d_in, d_sh, sb_pos, sb_v, best, depth, bb); });
  check("k_probe_diag", best);
  {   // packed sort: one u32 per entry between the passes; sa_v / best reused as its buffers
    uint32_t* pk = sa_v; uint32_t* ppos = sa_pos;
    timed("packed sort, both passes", [&] {
      k_sort_hist<0><<<gs, 256>>>(d_in, nullptr, d_sh, hist, bb, 0, SORT_LOW_BITS);
      k_sort_scan<<<n_shard, 1024>>>(d_sh, hist, nb0);
      k_sort_scatter_packed<0><<<gs, 256>>>(d_in, nullptr, d_sh, hist, pk, bb, SORT_LOW_BITS);
      k_sort_hist_packed<<<gs, 256>>>(pk, d_sh, hist, bb - SORT_LOW_BITS);
      k_sort_scan<<<n_shard, 1024>>>(d_sh, hist, nb1);
      k_sort_scatter_packed<1><<<gs, 256>>>(d_in, pk, d_sh, hist, ppos, bb, bb - SORT_LOW_BITS);
    });
    unsigned long long s2[2]; const uint32_t* arr[2] = {sb_pos, ppos};
    for (int k = 0; k < 2; k++) { cudaMemset(d_sum, 0, 8); k_checksum<<<1024, 256>>>(arr[k], n - 3, d_sum); cudaMemcpy(&s2[k], d_sum, 8, cudaMemcpyDeviceToHost); }
    printf("    packed sort: %s\n", s2[0] == s2[1] ? "same order" : "DIFFERENT order");
    CKX(cudaMemset(best, 0xFF, n * 4));
    timed("k_probe_diag<packed>", [&] { k_probe_diag<true><<<gp, 256, sm>>>(d_in, d_sh, ppos, nullptr, best, depth, bb); });
    check("k_probe_diag<packed>", best);
  }
  run_variants(d_in, d_sh, sb_pos, sb_v, best, depth, bb, max_npos, n_shard, timed, check);
  CKX(cudaDeviceSynchronize());
  return 0;
}
