/*
 * qlinear_oracle.c -- CPU restatement of nano-vllm-hw3's quantized-linear hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under nano-vllm-hw3_b200/ may import, link or
 * execute this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs call it, and only as the checker / reported CPU baseline.
 *
 * Parity pinning: the reference ships no golden vectors.  This
 * restatement is pinned against outputs of the reference itself, generated in the build
 * container by tests/golden/make_golden.py (reference torch functions on CPU, reference
 * Triton kernels under TRITON_INTERPRET=1) and committed as tests/golden/ (npz files);
 * tests/test_oracle_golden.py replays them.
 *
 * Every function cites the reference lines it restates (paths relative to the reference
 * repository root, nanovllm/utils/quantization.py unless stated otherwise).
 *
 * Formats: fmt 0 = int8 (qmax 127), fmt 1 = float8 e4m3fn (qmax 448).
 * bf16 values travel as uint16_t bit patterns, e4m3 values as uint8_t bit patterns.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_INT8 0
#define ORC_FP8 1

/* ---------------------------------------------------------------- scalar helpers */

static inline float bf16_to_f32(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

/* float -> bf16, round to nearest even (torch's .to(torch.bfloat16); Triton tl.cast). */
static inline uint16_t f32_to_bf16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x0040u); /* NaN */
  uint32_t lsb = (u >> 16) & 1u;
  u += 0x7fffu + lsb;
  return (uint16_t)(u >> 16);
}

/* e4m3fn bit pattern -> float (bias 7, no inf, 0x7f = NaN, subnormals 2^-9 .. 7*2^-9). */
static inline float e4m3_to_f32(uint8_t b) {
  int sign = b >> 7, e = (b >> 3) & 0xf, m = b & 7;
  float v;
  if (e == 0) v = ldexpf((float)m, -9);
  else if (e == 15 && m == 7) v = NAN;
  else v = ldexpf((float)(8 + m), e - 10);
  return sign ? -v : v;
}

/* float -> e4m3fn, round to nearest even, saturate to +-448 (cvt.rn.satfinite.e4m3x2.f32,
 * which is what Triton's tl.cast(.., tl.float8e4nv) lowers to, :39; torch's
 * .to(float8_e4m3fn) (:228) agrees for |v| <= 464, the only range the quantisers produce). */
static inline uint8_t f32_to_e4m3(float f) {
  uint8_t sign = signbit(f) ? 0x80 : 0;
  float a = fabsf(f);
  if (isnan(a)) return (uint8_t)(sign | 0x7f);
  if (a < 0.015625f) {                    /* below 2^-6: subnormal grid, step 2^-9 */
    int q = (int)rintf(a * 512.0f);       /* 0..8; 8 == smallest normal, same bits */
    return (uint8_t)(sign | q);
  }
  int e;
  (void)frexpf(a, &e);                    /* a = m * 2^e, m in [0.5,1) */
  e -= 1;                                 /* a = (1.x) * 2^e */
  int mant = (int)rintf(ldexpf(a, 3 - e)); /* 8..16, RNE (default rounding mode) */
  if (mant == 16) { mant = 8; e += 1; }
  int ef = e + 7;
  if (ef > 15 || (ef == 15 && mant - 8 > 6)) return (uint8_t)(sign | 0x7e); /* 448 */
  return (uint8_t)(sign | (ef << 3) | (mant - 8));
}

static inline float qmax_of(int fmt) { return fmt == ORC_FP8 ? 448.0f : 127.0f; }

/* ---------------------------------------------------------------- quantisers */

/*
 * Per-row (per-token) symmetric quantiser: restates the Triton kernel
 * _per_row_quant_kernel, quantization.py:26-41, called from per_row_quant :147-163.
 *   row_max   = max(max_k |x|, 1e-8)                    :27
 *   scale     = row_max / qmax ; scale_inv = qmax/row_max :29-34   (IEEE fp32 division)
 *   int8: y = (int8) round_half_away(x * scale_inv)     :36,41 (libdevice.round == roundf)
 *   fp8 : y = e4m3_rn(x * scale_inv)                    :36,39
 * x is bf16 [M,K] with row stride ldx (elements); q is [M,K] dense; s is [M].
 */
void orc_quant_rows(const uint16_t* x, int64_t M, int64_t K, int64_t ldx, int fmt,
                    uint8_t* q, float* s) {
  const float qmax = qmax_of(fmt);
#pragma omp parallel for schedule(static)
  for (int64_t m = 0; m < M; ++m) {
    const uint16_t* xr = x + m * ldx;
    float amax = 0.0f;
    for (int64_t k = 0; k < K; ++k) {
      float a = fabsf(bf16_to_f32(xr[k]));
      if (a > amax) amax = a;
    }
    amax = fmaxf(amax, 1e-8f);
    s[m] = amax / qmax;
    const float inv = qmax / amax;
    uint8_t* qr = q + m * K;
    for (int64_t k = 0; k < K; ++k) {
      float v = bf16_to_f32(xr[k]) * inv;
      if (fmt == ORC_FP8) qr[k] = f32_to_e4m3(v);
      else qr[k] = (uint8_t)(int8_t)roundf(v);
    }
  }
}

/*
 * Per-group weight quantiser: restates per_group_quant, quantization.py:194-234 (torch).
 *   amax   = max(|group|).clamp(min=1e-8)               :220
 *   scales = amax / max_val                             :221
 *   int8: q = clamp(round_half_even(x / scales), -128, 127)  :224-226 (torch.round == rintf)
 *   fp8 : q = e4m3(x / scales)                          :228
 * w is bf16 [N,K] dense; q [N,K]; s [N, K/g].  group_size == K gives the per-row layout.
 */
void orc_quant_groups(const uint16_t* w, int64_t N, int64_t K, int64_t g, int fmt,
                      uint8_t* q, float* s) {
  const float qmax = qmax_of(fmt);
  const int64_t G = K / g;
#pragma omp parallel for schedule(static)
  for (int64_t n = 0; n < N; ++n) {
    for (int64_t j = 0; j < G; ++j) {
      const uint16_t* wr = w + n * K + j * g;
      uint8_t* qr = q + n * K + j * g;
      float amax = 0.0f;
      for (int64_t k = 0; k < g; ++k) {
        float a = fabsf(bf16_to_f32(wr[k]));
        if (a > amax) amax = a;
      }
      amax = fmaxf(amax, 1e-8f);
      const float sc = amax / qmax;
      s[n * G + j] = sc;
      for (int64_t k = 0; k < g; ++k) {
        float v = bf16_to_f32(wr[k]) / sc;
        if (fmt == ORC_FP8) qr[k] = f32_to_e4m3(v);
        else {
          float r = rintf(v);
          r = fminf(fmaxf(r, -128.0f), 127.0f);
          qr[k] = (uint8_t)(int8_t)r;
        }
      }
    }
  }
}

/*
 * Per-tensor quantiser: restates per_tensor_quant, quantization.py:166-191.  The reference
 * keeps `scale` in the input dtype (bf16 0-dim tensor, :182-183) and divides in that dtype
 * (:187,:189): x/scale is a bf16 division (fp32 divide, rounded to bf16), then round/cast.
 * Returns the scale as the bf16 bit pattern widened to float in *scale_out.
 */
void orc_quant_tensor(const uint16_t* x, int64_t numel, int fmt, uint8_t* q, float* scale_out) {
  const float qmax = qmax_of(fmt);
  float amax = 0.0f;
  for (int64_t i = 0; i < numel; ++i) {
    float a = fabsf(bf16_to_f32(x[i]));
    if (a > amax) amax = a;
  }
  /* amax is a bf16 value; clamp(min=1e-8) happens in bf16, then bf16 / python scalar -> bf16 */
  const float amax_c = fmaxf(amax, bf16_to_f32(f32_to_bf16(1e-8f)));
  const float sc = bf16_to_f32(f32_to_bf16(amax_c / qmax));
  *scale_out = sc;
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < numel; ++i) {
    float v = bf16_to_f32(f32_to_bf16(bf16_to_f32(x[i]) / sc));
    if (fmt == ORC_FP8) q[i] = f32_to_e4m3(v);
    else {
      float r = rintf(v);
      r = fminf(fmaxf(r, -128.0f), 127.0f);
      q[i] = (uint8_t)(int8_t)r;
    }
  }
}

/* ---------------------------------------------------------------- GEMMs */

/* Exact int32 accumulators acc[m,n] = sum_k a[m,k]*b[n,k]: the tl.dot int8 path,
 * quantization.py:96,105-106. */
void orc_gemm_i8_acc(const int8_t* a, const int8_t* b, int64_t M, int64_t N, int64_t K,
                     int32_t* acc) {
#pragma omp parallel for schedule(static) collapse(2)
  for (int64_t m = 0; m < M; ++m)
    for (int64_t n = 0; n < N; ++n) {
      const int8_t* ar = a + m * K;
      const int8_t* br = b + n * K;
      int32_t s = 0;
      for (int64_t k = 0; k < K; ++k) s += (int32_t)ar[k] * (int32_t)br[k];
      acc[m * N + n] = s;
    }
}

/*
 * Per-row scaled W8A8 GEMM: restates _per_row_post_process_mm, quantization.py:54-144.
 *   acc (int32 exact, or fp32 for fp8)                  :93-106
 *   c = float(acc) * sa[m] * sb[n]   (left to right)    :111,121
 *   c += bias[n]                                        :123-133  (*)
 *   y = bf16(c)                                         :144
 * (*) the reference adds bias[:, None] (a row-index broadcast, :133); this restatement and
 * the CUDA path add bias[n] per column -- the deliberate deviation SURVEY.md section 4 records.
 * FP8 accumulation order is unspecified in the reference (tl.dot); here it is the exact
 * sum (double accumulation) rounded once to fp32, and tests allow a stated tolerance.
 * bias: bf16 [N] or NULL.  ldy: row stride of y in elements.  y_f32 (nullable, dense [M,N])
 * receives c before the bf16 cast -- the golden test needs it because TRITON_INTERPRET
 * truncates fp32->bf16 where the GPU's cvt.rn.bf16.f32 rounds to nearest even.
 */
void orc_gemm_rows(const uint8_t* qx, const float* sx, const uint8_t* qw, const float* sw,
                   const uint16_t* bias, int64_t M, int64_t N, int64_t K, int fmt,
                   uint16_t* y, int64_t ldy, float* y_f32) {
  float lut[256];
  for (int i = 0; i < 256; ++i) lut[i] = e4m3_to_f32((uint8_t)i);
#pragma omp parallel for schedule(static) collapse(2)
  for (int64_t m = 0; m < M; ++m)
    for (int64_t n = 0; n < N; ++n) {
      const uint8_t* ar = qx + m * K;
      const uint8_t* br = qw + n * K;
      float accf;
      if (fmt == ORC_INT8) {
        int32_t s = 0;
        for (int64_t k = 0; k < K; ++k) s += (int32_t)(int8_t)ar[k] * (int32_t)(int8_t)br[k];
        accf = (float)s;
      } else {
        double s = 0.0;
        for (int64_t k = 0; k < K; ++k) s += (double)lut[ar[k]] * (double)lut[br[k]];
        accf = (float)s;
      }
      float c = accf * sx[m];
      c = c * sw[n];
      if (bias) c += bf16_to_f32(bias[n]);
      y[m * ldy + n] = f32_to_bf16(c);
      if (y_f32) y_f32[m * N + n] = c;
    }
}

/*
 * Per-group W8A8 GEMM -- the new engine's semantics (last paragraph):
 *   P_j[m,n] = sum_{k in group j} qx[m,k]*qw[n,k]        (int32 exact / fp32 for fp8)
 *   t = 0; for j in 0..G-1: t = fmaf(float(P_j), sw[n,j], t)
 *   y = bf16( t * sx[m] + bias[n] )
 * The reference's per-group path (per_group_matmul :265-307) keeps activations
 * unquantised; orc_dequant_matmul_groups below restates that one.  `gp` is the partial
 * granularity: the kernel applies scales per min(g,128)-wide K block, so a group wider
 * than 128 contributes g/128 consecutive fmaf terms with the same scale.
 */
void orc_gemm_groups(const uint8_t* qx, const float* sx, const uint8_t* qw, const float* sw,
                     const uint16_t* bias, int64_t M, int64_t N, int64_t K, int64_t g,
                     int64_t gp, int fmt, uint16_t* y, int64_t ldy) {
  float lut[256];
  for (int i = 0; i < 256; ++i) lut[i] = e4m3_to_f32((uint8_t)i);
  const int64_t G = K / g;
#pragma omp parallel for schedule(static) collapse(2)
  for (int64_t m = 0; m < M; ++m)
    for (int64_t n = 0; n < N; ++n) {
      const uint8_t* ar = qx + m * K;
      const uint8_t* br = qw + n * K;
      float t = 0.0f;
      for (int64_t k0 = 0; k0 < K; k0 += gp) {
        float pf;
        if (fmt == ORC_INT8) {
          int32_t s = 0;
          for (int64_t k = k0; k < k0 + gp; ++k)
            s += (int32_t)(int8_t)ar[k] * (int32_t)(int8_t)br[k];
          pf = (float)s;
        } else {
          double s = 0.0;
          for (int64_t k = k0; k < k0 + gp; ++k) s += (double)lut[ar[k]] * (double)lut[br[k]];
          pf = (float)s;
        }
        t = fmaf(pf, sw[n * G + k0 / g], t);
      }
      float c = t * sx[m];
      if (bias) c += bf16_to_f32(bias[n]);
      y[m * ldy + n] = f32_to_bf16(c);
    }
}

/*
 * Weight-only dequant + fp32 matmul: restates per_group_matmul, quantization.py:265-307
 * (and per_tensor_matmul :237-262 with G == 1 and a broadcast scale).
 *   W = float(qw) * sw[n,j]                              :299-300
 *   y = bf16( x.float() @ W.T )                          :302   (fp32 GEMM; order unspecified,
 *                                                         restated as exact sum rounded once)
 *   y = bf16( y + bias )                                 :304-305 (bf16 + bf16 add)
 */
void orc_dequant_matmul_groups(const uint16_t* x, const uint8_t* qw, const float* sw,
                               const uint16_t* bias, int64_t M, int64_t N, int64_t K, int64_t g,
                               int fmt, uint16_t* y) {
  float lut[256];
  for (int i = 0; i < 256; ++i) lut[i] = e4m3_to_f32((uint8_t)i);
  const int64_t G = K / g;
  float* W = (float*)malloc(sizeof(float) * (size_t)N * (size_t)K);
#pragma omp parallel for schedule(static)
  for (int64_t n = 0; n < N; ++n)
    for (int64_t k = 0; k < K; ++k) {
      uint8_t b = qw[n * K + k];
      float v = fmt == ORC_INT8 ? (float)(int8_t)b : lut[b];
      W[n * K + k] = v * sw[n * G + k / g];
    }
#pragma omp parallel for schedule(static) collapse(2)
  for (int64_t m = 0; m < M; ++m)
    for (int64_t n = 0; n < N; ++n) {
      double s = 0.0;
      for (int64_t k = 0; k < K; ++k) s += (double)bf16_to_f32(x[m * K + k]) * (double)W[n * K + k];
      uint16_t yb = f32_to_bf16((float)s);
      if (bias) yb = f32_to_bf16(bf16_to_f32(yb) + bf16_to_f32(bias[n]));
      y[m * N + n] = yb;
    }
  free(W);
}

/* Conversion helpers exported for the Python side of the oracle. */
void orc_e4m3_decode(const uint8_t* in, int64_t n, float* out) {
  for (int64_t i = 0; i < n; ++i) out[i] = e4m3_to_f32(in[i]);
}
void orc_e4m3_encode(const float* in, int64_t n, uint8_t* out) {
  for (int64_t i = 0; i < n; ++i) out[i] = f32_to_e4m3(in[i]);
}
void orc_bf16_encode(const float* in, int64_t n, uint16_t* out) {
  for (int64_t i = 0; i < n; ++i) out[i] = f32_to_bf16(in[i]);
}
int orc_version(void) { return 1; }
