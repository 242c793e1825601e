// Stand-alone numerics + timing harness for the native kernels (no PyTorch needed):
//   nvcc ... -o build/selftest csrc/selftest.cu build/libtgb200.a   ;  ./build/selftest [case]
// Inputs are small multiples of 1/4 so TF32 products and FP32 sums are exact: any mismatch is a
// layout / descriptor bug, not rounding.
#include <cuda_runtime.h>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <vector>
#include "../include/tgb200.h"

#define CK(x)                                                                          \
  do {                                                                                 \
    cudaError_t e_ = (x);                                                              \
    if (e_ != cudaSuccess) {                                                           \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(2);                                                                         \
    }                                                                                  \
  } while (0)

static uint32_t rng_state = 12345u;
static float rnd_q() {  // in {-2,...,2}/4
  rng_state = rng_state * 1664525u + 1013904223u;
  return (static_cast<int>((rng_state >> 16) % 5) - 2) * 0.25f;
}

struct ConvCase {
  const char* name;
  int N, H, W, C, K, R, S, ph, pw, stats_mode;
};

static void conv_ref(const std::vector<float>& x, const std::vector<float>& w,
                     std::vector<float>& y, const ConvCase& c, int Ho, int Wo) {
  for (int n = 0; n < c.N; ++n)
    for (int ho = 0; ho < Ho; ++ho)
      for (int wo = 0; wo < Wo; ++wo)
        for (int k = 0; k < c.K; ++k) {
          float acc = 0.f;
          for (int r = 0; r < c.R; ++r) {
            int hi = ho + r - c.ph;
            if (hi < 0 || hi >= c.H) continue;
            for (int s = 0; s < c.S; ++s) {
              int wi = wo + s - c.pw;
              if (wi < 0 || wi >= c.W) continue;
              const float* xp = &x[((size_t)(n * c.H + hi) * c.W + wi) * c.C];
              const float* wp = &w[((size_t)(k * c.R + r) * c.S + s) * c.C];
              for (int ch = 0; ch < c.C; ++ch) acc += xp[ch] * wp[ch];
            }
          }
          y[((size_t)(n * Ho + ho) * Wo + wo) * c.K + k] = acc;
        }
}

static int run_conv_case(const ConvCase& c, bool check, int timing_iters) {
  const int Ho = c.H + 2 * c.ph - c.R + 1, Wo = c.W + 2 * c.pw - c.S + 1;
  const size_t nx = (size_t)c.N * c.H * c.W * c.C, nw = (size_t)c.K * c.R * c.S * c.C;
  const size_t ny = (size_t)c.N * Ho * Wo * c.K;
  std::vector<float> hx(nx), hw(nw), hy(ny), ref(ny);
  for (auto& v : hx) v = rnd_q();
  for (auto& v : hw) v = rnd_q();
  float *dx, *dw, *dy, *dstats;
  const size_t nstats = (size_t)c.N * c.K * 2;
  CK(cudaMalloc(&dx, nx * 4));
  CK(cudaMalloc(&dw, nw * 4));
  CK(cudaMalloc(&dy, ny * 4));
  CK(cudaMalloc(&dstats, nstats * 4));
  CK(cudaMemcpy(dx, hx.data(), nx * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dw, hw.data(), nw * 4, cudaMemcpyHostToDevice));
  CK(cudaMemset(dy, 0xFF, ny * 4));
  CK(cudaMemset(dstats, 0, nstats * 4));
  int rc = tgb_conv2d_fprop_tf32(dx, dw, dy, c.stats_mode ? dstats : nullptr, c.N, c.H, c.W, c.C,
                                 c.K, c.R, c.S, c.ph, c.pw, c.stats_mode, 0);
  if (rc) {
    printf("[%s] launch failed: %s\n", c.name, tgb_last_error());
    return 1;
  }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("[%s] kernel fault: %s\n", c.name, cudaGetErrorString(e));
    exit(3);
  }
  int bad = 0;
  if (check) {
    CK(cudaMemcpy(hy.data(), dy, ny * 4, cudaMemcpyDeviceToHost));
    conv_ref(hx, hw, ref, c, Ho, Wo);
    double maxerr = 0;
    for (size_t i = 0; i < ny; ++i) {
      double d = std::fabs((double)hy[i] - ref[i]);
      if (!(d <= 1e-3)) {
        if (bad < 12) {
          size_t k = i % c.K, p = i / c.K;
          size_t wo = p % Wo, ho = (p / Wo) % Ho, n = p / ((size_t)Wo * Ho);
          printf("  mismatch n=%zu ho=%zu wo=%zu k=%zu got=%g want=%g\n", n, ho, wo, k, hy[i],
                 ref[i]);
        }
        ++bad;
      }
      if (d > maxerr) maxerr = d;
    }
    if (c.stats_mode) {
      std::vector<float> hs(nstats);
      CK(cudaMemcpy(hs.data(), dstats, nstats * 4, cudaMemcpyDeviceToHost));
      int sbad = 0;
      const int groups = c.stats_mode == 1 ? c.N : 1;
      for (int gi = 0; gi < groups; ++gi)
        for (int k = 0; k < c.K; ++k) {
          double s1 = 0, s2 = 0;
          for (int n = 0; n < c.N; ++n) {
            if (c.stats_mode == 1 && n != gi) continue;
            for (int p = 0; p < Ho * Wo; ++p) {
              double v = ref[((size_t)n * Ho * Wo + p) * c.K + k];
              s1 += v;
              s2 += v * v;
            }
          }
          float g1 = hs[((size_t)gi * c.K + k) * 2], g2 = hs[((size_t)gi * c.K + k) * 2 + 1];
          if (std::fabs(g1 - s1) > 1e-3 * (1 + std::fabs(s1)) ||
              std::fabs(g2 - s2) > 1e-3 * (1 + std::fabs(s2))) {
            if (sbad < 6)
              printf("  stats mismatch g=%d k=%d got=(%g,%g) want=(%g,%g)\n", gi, k, g1, g2, s1, s2);
            ++sbad;
          }
        }
      bad += sbad;
    }
    printf("[%s] N=%d HxW=%dx%d C=%d K=%d %dx%d: max|err|=%.3g mismatches=%d -> %s\n", c.name, c.N,
           c.H, c.W, c.C, c.K, c.R, c.S, maxerr, bad, bad ? "FAIL" : "ok");
  }
  if (timing_iters > 0) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    float* st = c.stats_mode ? dstats : nullptr;
    for (int i = 0; i < 3; ++i)
      tgb_conv2d_fprop_tf32(dx, dw, dy, st, c.N, c.H, c.W, c.C, c.K, c.R, c.S, c.ph, c.pw,
                            c.stats_mode, 0);
    CK(cudaEventRecord(a));
    for (int i = 0; i < timing_iters; ++i)
      tgb_conv2d_fprop_tf32(dx, dw, dy, st, c.N, c.H, c.W, c.C, c.K, c.R, c.S, c.ph, c.pw,
                            c.stats_mode, 0);
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    ms /= timing_iters;
    double flop = 2.0 * c.N * Ho * Wo * (double)c.K * c.C * c.R * c.S;
    printf("[%s stats=%d] %.3f ms  %.1f TFLOP/s (tf32)\n", c.name, c.stats_mode, ms,
           flop / ms * 1e-9);
  }
  cudaFree(dx);
  cudaFree(dw);
  cudaFree(dy);
  cudaFree(dstats);
  return bad != 0;
}

static int run_wgrad_case(const ConvCase& c, bool check, int timing_iters) {
  const int Ho = c.H + 2 * c.ph - c.R + 1, Wo = c.W + 2 * c.pw - c.S + 1;
  const size_t nx = (size_t)c.N * c.H * c.W * c.C, nw = (size_t)c.K * c.R * c.S * c.C;
  const size_t ny = (size_t)c.N * Ho * Wo * c.K;
  std::vector<float> hx(nx), hdy(ny), hw(nw), ref(nw, 0.f);
  for (auto& v : hx) v = rnd_q();
  for (auto& v : hdy) v = rnd_q();
  float *dx, *ddy, *dw, *scratch = nullptr;
  long long sb = tgb_conv2d_wgrad_scratch_bytes(c.N, c.H, c.W, c.C, c.K, c.R, c.S, c.ph, c.pw);
  if (sb < 0) {
    printf("[wgrad %s] plan failed: %s\n", c.name, tgb_last_error());
    return 1;
  }
  CK(cudaMalloc(&dx, nx * 4));
  CK(cudaMalloc(&ddy, ny * 4));
  CK(cudaMalloc(&dw, nw * 4));
  if (sb) CK(cudaMalloc(&scratch, sb));
  CK(cudaMemcpy(dx, hx.data(), nx * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(ddy, hdy.data(), ny * 4, cudaMemcpyHostToDevice));
  CK(cudaMemset(dw, 0xFF, nw * 4));
  int rc = tgb_conv2d_wgrad_tf32(ddy, dx, dw, scratch, sb, c.N, c.H, c.W, c.C, c.K, c.R, c.S,
                                 c.ph, c.pw, 0, 0);
  if (rc) {
    printf("[wgrad %s] launch failed: %s\n", c.name, tgb_last_error());
    return 1;
  }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("[wgrad %s] kernel fault: %s\n", c.name, cudaGetErrorString(e));
    exit(3);
  }
  int bad = 0;
  if (check) {
    CK(cudaMemcpy(hw.data(), dw, nw * 4, cudaMemcpyDeviceToHost));
    for (int n = 0; n < c.N; ++n)
      for (int ho = 0; ho < Ho; ++ho)
        for (int wo = 0; wo < Wo; ++wo)
          for (int r = 0; r < c.R; ++r) {
            int hi = ho + r - c.ph;
            if (hi < 0 || hi >= c.H) continue;
            for (int s2 = 0; s2 < c.S; ++s2) {
              int wi = wo + s2 - c.pw;
              if (wi < 0 || wi >= c.W) continue;
              const float* xp = &hx[((size_t)(n * c.H + hi) * c.W + wi) * c.C];
              const float* gp = &hdy[((size_t)(n * Ho + ho) * Wo + wo) * c.K];
              for (int k = 0; k < c.K; ++k) {
                float gv = gp[k];
                if (gv == 0.f) continue;
                float* wp = &ref[((size_t)(k * c.R + r) * c.S + s2) * c.C];
                for (int ch = 0; ch < c.C; ++ch) wp[ch] += gv * xp[ch];
              }
            }
          }
    double maxerr = 0;
    for (size_t i = 0; i < nw; ++i) {
      double d = std::fabs((double)hw[i] - ref[i]);
      if (!(d <= 1e-3)) {
        if (bad < 12) {
          size_t ch = i % c.C, t = (i / c.C) % (c.R * c.S), k = i / ((size_t)c.C * c.R * c.S);
          printf("  wgrad mismatch k=%zu tap=%zu c=%zu got=%g want=%g\n", k, t, ch, hw[i], ref[i]);
        }
        ++bad;
      }
      if (d > maxerr) maxerr = d;
    }
    printf("[wgrad %s] N=%d HxW=%dx%d C=%d K=%d %dx%d scratch=%lld: max|err|=%.3g mismatches=%d -> %s\n",
           c.name, c.N, c.H, c.W, c.C, c.K, c.R, c.S, sb, maxerr, bad, bad ? "FAIL" : "ok");
  }
  if (timing_iters > 0) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    for (int i = 0; i < 3; ++i)
      tgb_conv2d_wgrad_tf32(ddy, dx, dw, scratch, sb, c.N, c.H, c.W, c.C, c.K, c.R, c.S, c.ph,
                            c.pw, 0, 0);
    CK(cudaEventRecord(a));
    for (int i = 0; i < timing_iters; ++i)
      tgb_conv2d_wgrad_tf32(ddy, dx, dw, scratch, sb, c.N, c.H, c.W, c.C, c.K, c.R, c.S, c.ph,
                            c.pw, 0, 0);
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    ms /= timing_iters;
    double flop = 2.0 * c.N * Ho * Wo * (double)c.K * c.C * c.R * c.S;
    printf("[wgrad %s] %.3f ms  %.1f TFLOP/s (tf32)\n", c.name, ms, flop / ms * 1e-9);
  }
  cudaFree(dx);
  cudaFree(ddy);
  cudaFree(dw);
  if (scratch) cudaFree(scratch);
  return bad != 0;
}

int main(int argc, char** argv) {
  const char* only = argc > 1 ? argv[1] : nullptr;
  int fails = 0;
  if (only && !strcmp(only, "shift_probe")) {
    // Does the MMA read a swizzled tile correctly when its start address is moved by whole
    // 128-byte rows? 1x1 conv on a single 32-pixel line; only pixels [shift, shift+8*ks) count.
    const int W = 32, C = 32, K = 128;
    std::vector<float> hx((size_t)W * C), hdy((size_t)W * K), hw((size_t)K * C);
    for (auto& v : hx) v = rnd_q();
    for (auto& v : hdy) v = rnd_q();
    float *dx, *ddy, *dw;
    CK(cudaMalloc(&dx, hx.size() * 4));
    CK(cudaMalloc(&ddy, hdy.size() * 4));
    CK(cudaMalloc(&dw, hw.size() * 4));
    CK(cudaMemcpy(dx, hx.data(), hx.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ddy, hdy.data(), hdy.size() * 4, cudaMemcpyHostToDevice));
    for (int shift = 0; shift <= 5; ++shift)
      for (int baseoff = 0; baseoff < 8; ++baseoff) {
        if (baseoff != 0 && baseoff != (shift & 7) && baseoff != (shift & 3) &&
            baseoff != ((8 - shift) & 7) && baseoff != ((4 - shift) & 3))
          continue;
        const int ks = 3;
        char env[64];
        snprintf(env, sizeof env, "%d,%d,%d", shift, ks, baseoff);
        setenv("TGB_WGRAD_DBG", env, 1);
        CK(cudaMemset(dw, 0xFF, hw.size() * 4));
        if (tgb_conv2d_wgrad_tf32(ddy, dx, dw, nullptr, 0, 1, 1, W, C, K, 1, 1, 0, 0, 0, 0)) {
          printf("launch failed: %s\n", tgb_last_error());
          return 1;
        }
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
          printf("shift=%d baseoff=%d: fault %s\n", shift, baseoff, cudaGetErrorString(e));
          return 3;
        }
        CK(cudaMemcpy(hw.data(), dw, hw.size() * 4, cudaMemcpyDeviceToHost));
        int bad = 0;
        for (int k = 0; k < K; ++k)
          for (int c = 0; c < C; ++c) {
            float ref = 0.f;
            for (int p = shift; p < shift + 8 * ks; ++p) ref += hdy[(size_t)p * K + k] * hx[(size_t)p * C + c];
            if (std::fabs(hw[(size_t)k * C + c] - ref) > 1e-3) ++bad;
          }
        printf("shift_probe shift=%d base_offset=%d -> %s (%d mismatches)\n", shift, baseoff,
               bad ? "FAIL" : "ok", bad);
      }
    unsetenv("TGB_WGRAD_DBG");
    return 0;
  }
  if (only && !strcmp(only, "wgrad_sweep")) {
    // lbo, sbo, layout type, TMA swizzle mode
    const char* variants[] = {"4096,512,1,2", "512,4096,1,2", "4096,1024,1,2", "1024,4096,1,2",
                              "4096,256,1,2", "4096,512,2,1", "4096,1024,2,1", "1024,4096,2,1",
                              "4096,512,1,1", "4096,1024,2,2"};
    ConvCase cs[] = {{"sweep_k128c64", 1, 16, 16, 64, 128, 1, 1, 0, 0, 0},
                     {"sweep_k32c32", 2, 8, 16, 32, 32, 1, 1, 0, 0, 0}};
    for (auto v : variants) {
      setenv("TGB_WGRAD_DESC", v, 1);
      printf("== variant lbo,sbo,layout,tma = %s\n", v);
      for (auto& c : cs) run_wgrad_case(c, true, 0);
    }
    return 0;
  }
  ConvCase small[] = {
      {"gemm1x1", 2, 8, 16, 32, 32, 1, 1, 0, 0, 0},
      {"gemm1x1_k64c64", 1, 16, 16, 64, 64, 1, 1, 0, 0, 2},
      {"conv3x3_a", 2, 16, 16, 64, 64, 3, 3, 1, 1, 1},
      {"conv3x3_b", 3, 12, 12, 96, 128, 3, 3, 1, 1, 1},
      {"conv3x3_multiimg", 5, 6, 6, 32, 256, 3, 3, 1, 1, 1},
      {"conv3x3_partial", 2, 10, 20, 32, 32, 3, 3, 1, 1, 2},
      {"conv1x7", 2, 8, 24, 32, 64, 1, 7, 0, 3, 0},
      {"conv7x1", 2, 24, 8, 32, 64, 7, 1, 3, 0, 0},
      {"conv3x3_c512", 1, 8, 8, 512, 256, 3, 3, 1, 1, 0},
      {"conv3x3_c20", 1, 8, 8, 20, 32, 3, 3, 1, 1, 0},
      {"gemm1x1_k40", 2, 8, 8, 32, 40, 1, 1, 0, 0, 2},
      {"conv3x3_k72", 1, 8, 8, 64, 72, 3, 3, 1, 1, 1},
      {"linear_2048_1000", 40, 1, 1, 2048, 1000, 1, 1, 0, 0, 0},
      {"conv3x3_splitk_in", 2, 6, 6, 256, 256, 3, 3, 1, 1, 1},
      {"conv3x3_splitk_bn", 3, 12, 12, 512, 64, 3, 3, 1, 1, 2},
  };
  for (auto& c : small) {
    if (only && strcmp(only, c.name) && strcmp(only, "all")) continue;
    fails += run_conv_case(c, true, 0);
    fails += run_wgrad_case(c, true, 0);
  }
  // haloed-tile forward variant (8x16 pixel tiles, input read once per tile), forced on
  ConvCase halo_cases[] = {
      {"halo_64_64", 2, 40, 24, 64, 64, 3, 3, 1, 1, 1},
      {"halo_48_32", 3, 16, 20, 48, 32, 3, 3, 1, 1, 2},
      {"halo_128_32_pad0", 1, 34, 18, 128, 32, 3, 3, 0, 0, 0},
      {"halo_1x3", 2, 16, 16, 32, 64, 1, 3, 0, 1, 1},
      {"halo_3x1_k40", 2, 18, 9, 32, 40, 3, 1, 1, 0, 0},
      {"halo_dgrad_pad", 2, 20, 12, 64, 64, 3, 3, 2, 2, 0},
  };
  for (auto& c : halo_cases) {
    if (only && strcmp(only, c.name) && strcmp(only, "all") && strcmp(only, "halo")) continue;
    setenv("TGB_FPROP_HALO", "2", 1);
    fails += run_conv_case(c, true, 0);           // filter bank resident in smem where it fits
    setenv("TGB_FPROP_RESW", "0", 1);
    fails += run_conv_case(c, true, 0);           // weights streamed with every stage
    unsetenv("TGB_FPROP_RESW");
    unsetenv("TGB_FPROP_HALO");
  }
  ConvCase perf[] = {
      {"perf_64_192", 8, 192, 192, 64, 64, 3, 3, 1, 1, 0},
      {"perf_128_96", 16, 96, 96, 128, 128, 3, 3, 1, 1, 0},
      {"perf_256_48", 16, 48, 48, 256, 256, 3, 3, 1, 1, 0},
      {"perf_512_24", 40, 24, 24, 512, 512, 3, 3, 1, 1, 0},
      {"perf_1024_12", 40, 12, 12, 1024, 1024, 3, 3, 1, 1, 0},
      {"perf_1x1_256", 16, 56, 56, 256, 256, 1, 1, 0, 0, 0},
  };
  for (auto& c : perf) {
    if (only && strcmp(only, c.name) && strcmp(only, "all") && strcmp(only, "perf")) continue;
    if (!only) continue;
    fails += run_conv_case(c, false, 10);
    fails += run_wgrad_case(c, false, 10);
  }
  int unet_batch = 40;
  if (only && !strncmp(only, "unetb", 5)) {     // "unetb16": the layer table at micro-batch 16
    unet_batch = atoi(only + 5);
    only = "unet";
  }
  if (only && (!strcmp(only, "unet") || !strncmp(only, "unet_", 5))) {
    // every distinct 3x3 layer of U-Net (5,64) at the benchmark's micro-batch of 40
    struct { int C, K, HW; } layers[] = {
        {64, 64, 192}, {64, 128, 96}, {128, 128, 96}, {128, 256, 48}, {256, 256, 48},
        {256, 512, 24}, {512, 512, 24}, {512, 1024, 12}, {1024, 1024, 12}, {1024, 2048, 6},
        {2048, 2048, 6}, {2048, 1024, 6}, {2048, 512, 12}, {512, 512, 12}, {1024, 256, 24},
        {256, 256, 24}, {512, 128, 48}, {128, 128, 48}, {256, 64, 96}, {64, 64, 96},
        {128, 32, 192}, {32, 32, 192}};
    for (auto& l : layers) {
      char name[64];
      snprintf(name, sizeof name, "unet_%d_%d_%d", l.C, l.K, l.HW);
      if (strcmp(only, "unet") && strcmp(only, name)) continue;
      ConvCase c{name, unet_batch, l.HW, l.HW, l.C, l.K, 3, 3, 1, 1, 0};
      run_conv_case(c, false, 5);
      ConvCase cs{name, unet_batch, l.HW, l.HW, l.C, l.K, 3, 3, 1, 1, 1};
      run_conv_case(cs, false, 5);
      run_wgrad_case(c, false, 5);
    }
  }
  printf("selftest: %s\n", fails ? "FAILURES" : "all ok");
  return fails ? 1 : 0;
}
