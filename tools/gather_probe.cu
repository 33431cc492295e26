// tools/gather_probe.cu -- development probe for the gather kernel: runs
// row_copy_kernel on a synthetic slab with per-CTA timestamps so that the
// critical path (launch, id load, first piece latency, drain) can be read off.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/gather_probe \
//        tools/gather_probe.cu rela_b200/csrc/row_copy.cu
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include "../rela_b200/csrc/kernels.h"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

int main(int argc, char** argv) {
  int slots = argc > 1 ? atoi(argv[1]) : 131072;
  int B = argc > 2 ? atoi(argv[2]) : 512;
  int piece = argc > 3 ? atoi(argv[3]) : 8192;
  int ctas_per_sm = argc > 4 ? atoi(argv[4]) : 1;
  int max_stages = argc > 5 ? atoi(argv[5]) : 32;
  const uint32_t row = 28224;
  const int nf = 2;
  uint8_t* slab[nf];
  uint8_t* out[8][nf];
  for (int f = 0; f < nf; ++f) {
    CK(cudaMalloc(&slab[f], (size_t)slots * row));
    CK(cudaMemset(slab[f], f + 1, (size_t)slots * row));
    for (int k = 0; k < 8; ++k) CK(cudaMalloc(&out[k][f], (size_t)B * row));
  }
  std::mt19937 g(1);
  const int NID = 16;
  int32_t* ids[NID];
  for (int k = 0; k < NID; ++k) {
    std::vector<int32_t> h(B);
    for (auto& v : h) v = (int32_t)(g() % (uint32_t)slots);
    std::sort(h.begin(), h.end());
    CK(cudaMalloc(&ids[k], B * 4));
    CK(cudaMemcpy(ids[k], h.data(), B * 4, cudaMemcpyHostToDevice));
  }
  rr::CopyParams p;
  memset(&p, 0, sizeof(p));
  uint32_t pieces = (row + piece - 1) / piece;
  uint32_t pb = (((row + pieces - 1) / pieces) + 15u) & ~15u;
  uint32_t items = 0;
  for (int f = 0; f < nf; ++f) {
    rr::CopyField& c = p.bulk[f];
    c.src = slab[f];
    c.src_row_stride = row;
    c.dst_row_stride = row;
    c.dst_chunk_stride = (uint64_t)B * row;
    c.chunk_bytes = row;
    c.chunks = 1;
    c.pieces = pieces;
    c.piece_bytes = pb;
    c.begin = items;
    items += (uint32_t)B * pieces;
  }
  p.n_bulk = nf;
  p.bulk_items = items;
  p.rows = B;
  p.slots = slots;
  p.stage_bytes = (pb + 127u) & ~127u;
  const int sms = 148;
  rr::CopyLaunch cfg;
  cfg.block = 128;
  cfg.grid = std::min<int>(sms * ctas_per_sm, (int)items);
  size_t hdr = rr::kCopyHeaderBytes;
  size_t lim = 227 * 1024 / ctas_per_sm - 1024;
  int per_cta = (items + cfg.grid - 1) / cfg.grid;
  int fit = (int)((lim - hdr) / p.stage_bytes);
  p.stages = std::max(1, std::min({per_cta, fit, max_stages, rr::kMaxCopyStages}));
  cfg.smem = hdr + (size_t)p.stages * p.stage_bytes;
  unsigned long long* probe;
  CK(cudaMalloc(&probe, (size_t)cfg.grid * 64));
  cudaStream_t st;
  CK(cudaStreamCreate(&st));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  printf("slots %d B %d piece %u x%u items %u grid %d stages %d smem %zu\n", slots, B, pb, pieces, items, cfg.grid,
         p.stages, cfg.smem);
  auto run = [&](int k, bool with_probe) {
    for (int f = 0; f < nf; ++f) p.bulk[f].dst = out[k % 8][f];
    p.ids = ids[k % NID];
    p.probe = with_probe ? probe : nullptr;
    CK(rr::launch_row_copy(p, cfg, st));
  };
  for (int k = 0; k < 20; ++k) run(k, false);
  CK(cudaStreamSynchronize(st));
  // (a) isolated launches, each bracketed by events
  float tot = 0;
  for (int k = 0; k < 64; ++k) {
    CK(cudaEventRecord(e0, st));
    run(k, false);
    CK(cudaEventRecord(e1, st));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    tot += ms;
  }
  printf("event-bracketed single launch: %.2f us\n", tot / 64 * 1e3);
  // (b) back-to-back launches
  CK(cudaEventRecord(e0, st));
  for (int k = 0; k < 256; ++k) run(k, false);
  CK(cudaEventRecord(e1, st));
  CK(cudaEventSynchronize(e1));
  float ms;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  printf("back-to-back: %.2f us per launch (%.0f GB/s algorithmic)\n", ms / 256 * 1e3,
         2.0 * nf * row * B / (ms / 256 * 1e-3) / 1e9);
  // (c) one probed launch
  run(3, true);
  CK(cudaStreamSynchronize(st));
  std::vector<unsigned long long> h((size_t)cfg.grid * 8);
  CK(cudaMemcpy(h.data(), probe, h.size() * 8, cudaMemcpyDeviceToHost));
  unsigned long long g0 = ~0ull, g1 = 0;
  double a[5] = {0, 0, 0, 0, 0}, mx[5] = {0, 0, 0, 0, 0};
  unsigned long long gs_max = 0;
  for (int c = 0; c < cfg.grid; ++c) {
    unsigned long long* r = &h[(size_t)c * 8];
    g0 = std::min(g0, r[0]);
    g1 = std::max(g1, r[7]);
    gs_max = std::max(gs_max, r[0]);
    double v[5] = {(double)(r[2] - r[1]), (double)(r[3] - r[2]), (double)(r[4] - r[3]), (double)(r[5] - r[4]),
                   (double)(r[6] - r[5])};
    for (int i = 0; i < 5; ++i) {
      a[i] += v[i] / cfg.grid;
      mx[i] = std::max(mx[i], v[i]);
    }
  }
  printf("probe: kernel span (globaltimer) %.2f us; last CTA started %.2f us after the first\n", (g1 - g0) / 1e3,
         (gs_max - g0) / 1e3);
  const char* names[5] = {"start->pdl_wait passed", "ids + issue first load", "first load latency", "remaining items loop",
                          "drain stores"};
  for (int i = 0; i < 5; ++i) printf("  %-26s avg %8.0f cyc  max %8.0f cyc\n", names[i], a[i], mx[i]);
  return 0;
}
