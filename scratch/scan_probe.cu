// Phase timeline of tile_scan_kernel (clock64 stamps by thread 0).  Build: see scripts/gpu_scan_probe.sh
#define SPLAT_SCAN_PROBE 1
#include "../splatam_b200/csrc/binning.cu"
#include <cstdio>
#include <vector>
#include <algorithm>
namespace splat { int g_prof_on = 0; void prof_record(int, cudaStream_t, bool) {}
  int g_launch_count = 0; int check_cuda(cudaError_t e, const char*) { return e != cudaSuccess; } }
int main() {
  using namespace splat;
  const int T = 75 * 43;   // 1200x680 / 16x16
  std::vector<uint32_t> h((size_t)T * CSTRIDE, 0u);
  srand(1);
  for (int t = 0; t < T; ++t) h[(size_t)t * CSTRIDE] = 200 + rand() % 1500;
  uint32_t *cnt, *cur, *order, *nr_dev, *nr_host; uint2* ranges; ImageHeader* hdr;
  cudaMalloc(&cnt, h.size() * 4); cudaMalloc(&cur, h.size() * 4); cudaMalloc(&order, T * 4); cudaMalloc(&ranges, T * 8);
  cudaMalloc(&hdr, 256); cudaMalloc(&nr_dev, 4);
  cudaHostAlloc(&nr_host, 4, cudaHostAllocMapped);
  cudaMemcpy(cnt, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int variant = 0; variant < 2; ++variant) {
    uint32_t* out = variant ? nr_host : nr_dev;
    for (int rep = 0; rep < 4; ++rep) {
      cudaEventRecord(e0);
      tile_scan_kernel<<<SCAN_CTAS, SCAN_THREADS>>>(T, cnt, cur, ranges, hdr, out, 2048u, order);
      cudaEventRecord(e1);
      cudaDeviceSynchronize();
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      long long s[16]; cudaMemcpyFromSymbol(s, g_scan_stamps, sizeof(s));
      printf("%s rep %d: %.2f us |", variant ? "host-mapped" : "device", rep, ms * 1e3f);
      for (int i = 1; i < 7; ++i) printf(" %d:%lld", i, s[i] - s[i - 1]);
      printf("  total %lld cycles  err=%d\n", s[6] - s[0], (int)cudaGetLastError());
    }
  }
  // check against a host scan
  std::vector<uint2> hr(T); std::vector<uint32_t> ho(T), hc(h.size());
  cudaMemcpy(hr.data(), ranges, T * 8, cudaMemcpyDeviceToHost); cudaMemcpy(ho.data(), order, T * 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(hc.data(), cur, h.size() * 4, cudaMemcpyDeviceToHost);
  uint32_t run = 0, mx = 0; int bad = 0;
  for (int t = 0; t < T; ++t) { uint32_t c = h[(size_t)t * CSTRIDE]; mx = c > mx ? c : mx;
    if (hr[t].x != run || hr[t].y != run + c || hc[(size_t)t * CSTRIDE] != run) ++bad; run += c; }
  std::vector<int> seen(T, 0); uint32_t prev_cls = 0;
  for (int k = 0; k < T; ++k) { uint32_t t = ho[k]; if (t >= (uint32_t)T || seen[t]++) { ++bad; continue; }
    uint32_t cls = 63u - std::min(63u, (uint32_t)((float)h[(size_t)t * CSTRIDE] * (64.f / ((float)mx + 1.f))));
    if (cls < prev_cls) ++bad; prev_cls = cls; }
  printf("R host %u device %u  bad=%d\n", run, *nr_host, bad);
  return bad != 0;
}
