// ref_bdd_wrap.cc -- C entry point around the reference's OWN calculate_block_density_dist, compiled from
// /root/reference/block_density_dist.cpp where it lies (main renamed, never called).  TEST INFRASTRUCTURE.
#define main ref_block_density_dist_main
#include "block_density_dist.cpp"   // found through -I$(REF): calculate_block_density_dist :47-86
#undef main

extern "C" void ref_block_density_dist(int n, const int* rowptr, const int* colind, int bsize, long long* dist10) {
  std::vector<std::vector<int>> e(n);
  for (int i = 0; i < n; ++i) e[i].assign(colind + rowptr[i], colind + rowptr[i + 1]);
  std::array<ll, 10> d = calculate_block_density_dist(e, n, bsize);
  for (int i = 0; i < 10; ++i) dist10[i] = d[i];
}
