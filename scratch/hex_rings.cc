// Histogram of flood depth (rings per placement) in random hex 11x11 playouts (host-side experiment).
#include <cstdio>
#include <cstdlib>
#include <random>
#include "../open_spiel-1_b200/csrc/bsg_games.cuh"
using namespace bsg;
typedef Hex H;
int main() {
  H::Tables* t = new H::Tables(); H::fill_tables(*t);
  std::mt19937 rng(1);
  long hist[64] = {0}; long steps = 0, floods = 0;
  long warpmax_sum = 0, warp_steps = 0;
  const int W = 32;
  for (int rep = 0; rep < 200; ++rep) {
    H::State s[W]; bool live[W];
    for (int l = 0; l < W; ++l) { H::init(s[l]); live[l] = true; }
    for (int step = 0; step < 121; ++step) {
      int mx = 0; bool any = false;
      for (int l = 0; l < W; ++l) {
        if (!live[l]) continue; any = true;
        // replicate step with ring counting
        H::State& S = s[l];
        int legal[128]; int n = H::legal(S, legal);
        int cell = legal[rng() % n];
        // count rings: copy of place logic
        u32 nb[4]; for (int w = 0; w < 4; ++w) nb[w] = t->nb[cell][w];
        u32 flags = nb[3] >> (28 + 2 * S.mv);
        u32 hitA = 0, hitB = 0;
        for (int w = 0; w < 4; ++w) { hitA |= nb[w] & S.cA[w]; hitB |= nb[w] & S.cB[w]; }
        bool tA = hitA || (flags & 1), tB = hitB || (flags & 2);
        int rings = 0;
        if ((tA || tB) && !(tA && tB)) {
          u32 P[4], fr[4], anyf = 0;
          for (int w = 0; w < 4; ++w) { P[w] = S.cP[w]; fr[w] = nb[w] & P[w]; anyf |= fr[w]; }
          while (anyf) {
            ++rings;
            u32 d[4];
            for (int w = 0; w < 4; ++w) P[w] &= ~fr[w];
            H::dilate(fr, *t, d);
            anyf = 0;
            for (int w = 0; w < 4; ++w) { fr[w] = d[w] & P[w]; anyf |= fr[w]; }
          }
        }
        hist[rings]++; steps++; if (rings) floods++;
        if (rings > mx) mx = rings;
        H::apply(S, cell, *t);
        if (H::terminal(S)) live[l] = false;
      }
      if (any) { warpmax_sum += mx; warp_steps++; }
    }
  }
  printf("steps %ld floods %ld (%.1f%%)\n", steps, floods, 100.0 * floods / steps);
  for (int i = 0; i < 20; ++i) if (hist[i]) printf("rings %2d: %7ld (%.2f%%)\n", i, hist[i], 100.0 * hist[i] / steps);
  printf("mean over warp-steps of max rings: %.2f\n", (double)warpmax_sum / warp_steps);
  return 0;
}
