// Histogram of Newton iterations per steering call on field-scale pairs (host build of csrc/krrt_math.cuh).
// Diagnostic for kernel tail latency: a warp runs as long as its slowest lane.
#include <cstdio>
#include <random>
static int g_it[3];
#define KRRT_ITER_HOOK(which) ++g_it[which];
#include "../../robosoccer_motion_planning_b200/csrc/krrt_math.cuh"
int main(int argc, char** argv) {
  int n = argc > 1 ? atoi(argv[1]) : 1000000;
  double radius = argc > 2 ? atof(argv[2]) : 0.0;  // > 0: second state within this 4-D distance of the first
  std::mt19937_64 rng(1);
  std::uniform_real_distribution<double> ux(-9, 9), uy(-6, 6), uv(-2, 2), u(-1, 1);
  long hist[512] = {0};
  long tot = 0; int mx = 0; long three = 0; long per[3] = {0,0,0};
  for (int i = 0; i < n; ++i) {
    krrt::State4 a = {ux(rng), uy(rng), uv(rng), uv(rng)}, b = {ux(rng), uy(rng), uv(rng), uv(rng)};
    if (radius > 0) { b.x = a.x + radius * u(rng); b.y = a.y + radius * u(rng); b.vx = a.vx + radius * u(rng); b.vy = a.vy + radius * u(rng); }
    g_it[0] = g_it[1] = g_it[2] = 0;
    double t, c;
    krrt::steer(a, b, 0.5, t, c);
    int k = g_it[0] + g_it[1] + g_it[2];
    if (g_it[1]) ++three;
    for (int q = 0; q < 3; ++q) per[q] += g_it[q];
    hist[k < 511 ? k : 511]++; tot += k; if (k > mx) mx = k;
  }
  printf("mean %.2f max %d  stationary-path calls %ld  per-loop means: root %.2f left %.2f right %.2f\n", (double)tot / n, mx, three, (double)per[0]/n, (double)per[1]/n, (double)per[2]/n);
  long acc = 0;
  for (int k = 0; k < 512; ++k) { acc += hist[k]; if (hist[k]) printf("%3d: %8ld  cum %.5f\n", k, hist[k], (double)acc / n); }
}
