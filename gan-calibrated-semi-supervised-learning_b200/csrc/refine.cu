// Refined-patch / letterbox generation on the device: crop a (cx,cy,w,h) box out of a device-resident uint8 RGB image,
// grey-pad (128) to a square, resize to S x S with PIL's BICUBIC (antialiased, 8-bit separable passes), ToTensor +
// Normalize(0.5, 0.5) -> fp32 NCHW in [-1, 1].  Replaces the CPU/PIL round trip of
//   cgan/cgan_train_enhanced.py:37-137  get_refined_patch_batch (mode 1: delta -> hard-clamp box update, the [0.1,0.9] /
//                                       [0.05,0.8] clamps, the "< 10 px -> fall back to pred_box" rule, int() truncation)
//   cgan/dataset.py:103-124             CalibratorDataset._letterbox (mode 0: plain box, int() truncation)
//   cgan/inference.py:51-66             crop_patch + letterbox (mode 2: PIL rounds float crop coordinates)
// The resize restates Pillow's ImagingResample (src/libImaging/Resample.c, BICUBIC a = -0.5, support 2 scaled by the
// down-scale factor, coefficients normalised in double and quantised to 22-bit fixed point, horizontal pass into uint8,
// then vertical pass), so the patches are BIT-IDENTICAL to the reference's - the double arithmetic below avoids FMA
// contraction (__dmul_rn / __dadd_rn) because Pillow's x86-64 builds do not contract.
#include "common.cuh"

namespace {
constexpr int RF_PRECISION_BITS = 32 - 8 - 2;

struct RefineGeom {  // per sample
  int img, ix1, iy1, cw, ch;  // image index, crop origin and size in pixels
  int padl, padt, Q;          // left / top grey padding and side of the padded square
  int ksize, valid;           // taps per output pixel; 0 = degenerate box (patch filled with zeros, like the reference's except branch)
};

__device__ __forceinline__ double bicubic_filter(double x) {
  const double a = -0.5;
  if (x < 0.0) x = -x;
  if (x < 1.0) return __dadd_rn(__dmul_rn(__dmul_rn(__dadd_rn(__dmul_rn(a + 2.0, x), -(a + 3.0)), x), x), 1.0);
  if (x < 2.0) return __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(x, -5.0), x), 8.0), x), -4.0), a);
  return 0.0;
}

__device__ __forceinline__ float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

// cgan/losses.py:108-150 with training=False (hard clamps), float32 like the reference's torch CPU ops
__device__ __forceinline__ void apply_delta_hard(const float* bb, const float* dl, float* out) {
  float dc[4];
  for (int i = 0; i < 4; ++i) dc[i] = clampf(dl[i], -1.5f, 1.5f);
  const float cx = bb[0] + dc[0] * bb[2], cy = bb[1] + dc[1] * bb[3];
  const float w = bb[2] * expf(clampf(dc[2], -1.f, 1.f)), h = bb[3] * expf(clampf(dc[3], -1.f, 1.f));
  out[0] = clampf(cx, 0.05f, 0.95f); out[1] = clampf(cy, 0.05f, 0.95f);
  out[2] = clampf(w, 0.02f, 0.8f); out[3] = clampf(h, 0.02f, 0.8f);
}

__device__ __forceinline__ int py_round(double v) {  // Python round(): half to even, as PIL's Image.crop applies to float boxes
  return (int)rint(v);
}

// One thread per sample: the crop rectangle.  box_out (nullable, [B][4]) receives the box that was cropped.
__global__ void refine_geom_kernel(const int* __restrict__ img_h, const int* __restrict__ img_w, const int* __restrict__ sample_img,
                                   const float* __restrict__ pred_box, const float* __restrict__ delta, int mode, int S,
                                   RefineGeom* __restrict__ geom, float* __restrict__ box_out, int B, int kmax) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  RefineGeom g;
  g.img = sample_img[b];
  const int W = img_w[g.img], H = img_h[g.img];
  float pb[4], rb[4];
  for (int i = 0; i < 4; ++i) pb[i] = pred_box[b * 4 + i];
  if (mode == 1 && delta) {
    float dl[4];
    for (int i = 0; i < 4; ++i) dl[i] = delta[b * 4 + i];
    apply_delta_hard(pb, dl, rb);
    rb[0] = clampf(rb[0], 0.1f, 0.9f); rb[1] = clampf(rb[1], 0.1f, 0.9f);
    rb[2] = clampf(rb[2], 0.05f, 0.8f); rb[3] = clampf(rb[3], 0.05f, 0.8f);
  } else {
    for (int i = 0; i < 4; ++i) rb[i] = pb[i];
  }
  auto corners = [&](const float* bx, double& x1, double& y1, double& x2, double& y2) {
    const double px = __dmul_rn((double)bx[0], (double)W), py = __dmul_rn((double)bx[1], (double)H);
    const double pw = __dmul_rn((double)bx[2], (double)W), ph = __dmul_rn((double)bx[3], (double)H);
    x1 = fmax(0.0, __dadd_rn(px, -pw / 2)); y1 = fmax(0.0, __dadd_rn(py, -ph / 2));
    x2 = fmin((double)W, __dadd_rn(px, pw / 2)); y2 = fmin((double)H, __dadd_rn(py, ph / 2));
  };
  double x1, y1, x2, y2;
  corners(rb, x1, y1, x2, y2);
  if (mode == 1 && (x2 <= x1 || y2 <= y1 || (x2 - x1) < 10 || (y2 - y1) < 10)) {  // fall back to the un-refined box
    corners(pb, x1, y1, x2, y2);
    for (int i = 0; i < 4; ++i) rb[i] = pb[i];
  }
  int ix1, iy1, ix2, iy2;
  if (mode == 2) { ix1 = py_round(x1); iy1 = py_round(y1); ix2 = py_round(x2); iy2 = py_round(y2); }
  else { ix1 = (int)x1; iy1 = (int)y1; ix2 = (int)x2; iy2 = (int)y2; }
  g.ix1 = ix1; g.iy1 = iy1; g.cw = ix2 - ix1; g.ch = iy2 - iy1;
  g.valid = g.cw > 0 && g.ch > 0;
  const int pad_w = max(g.ch - g.cw, 0), pad_h = max(g.cw - g.ch, 0);
  g.padl = pad_w / 2; g.padt = pad_h / 2;
  g.Q = max(g.cw, g.ch);
  // Pillow precompute_coeffs: filterscale = max(scale, 1), support = 2 * filterscale, ksize = ceil(support) * 2 + 1
  const double scale = (double)g.Q / (double)S;
  const double fs = scale < 1.0 ? 1.0 : scale;
  g.ksize = (int)ceil(__dmul_rn(2.0, fs)) * 2 + 1;
  if (g.ksize > kmax || g.Q <= 0) g.valid = 0;
  geom[b] = g;
  if (box_out)
    for (int i = 0; i < 4; ++i) box_out[b * 4 + i] = rb[i];
}

// One thread per (sample, output index): bounds + 22-bit fixed-point coefficients (identical for both axes: the padded
// crop is a square resized to a square).  ktab: [B][S][kmax] ints, bounds: [B][S][2] ints (xmin, count).
__global__ void refine_coeff_kernel(const RefineGeom* __restrict__ geom, int S, int kmax, int* __restrict__ ktab,
                                    int* __restrict__ bounds, int B) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * S) return;
  const int b = i / S, xx = i % S;
  const RefineGeom g = geom[b];
  int* k = ktab + (long long)i * kmax;
  if (!g.valid) { bounds[2 * i] = 0; bounds[2 * i + 1] = 0; return; }
  const double scale = (double)g.Q / (double)S;
  const double fs = scale < 1.0 ? 1.0 : scale;
  const double support = __dmul_rn(2.0, fs);
  const double center = __dmul_rn(__dadd_rn((double)xx, 0.5), scale);  // in0 = 0
  const double ss = 1.0 / fs;
  int xmin = (int)__dadd_rn(__dadd_rn(center, -support), 0.5);
  if (xmin < 0) xmin = 0;
  int xmax = (int)__dadd_rn(__dadd_rn(center, support), 0.5);
  if (xmax > g.Q) xmax = g.Q;
  xmax -= xmin;
  double ww = 0.0;
  for (int x = 0; x < xmax; ++x) {
    const double w = bicubic_filter(__dmul_rn(__dadd_rn(__dadd_rn((double)(x + xmin), -center), 0.5), ss));
    ww = __dadd_rn(ww, w);
  }
  for (int x = 0; x < xmax; ++x) {
    double w = bicubic_filter(__dmul_rn(__dadd_rn(__dadd_rn((double)(x + xmin), -center), 0.5), ss));
    if (ww != 0.0) w = w / ww;
    const double sc = __dmul_rn(w, (double)(1 << RF_PRECISION_BITS));
    k[x] = w < 0 ? (int)__dadd_rn(-0.5, sc) : (int)__dadd_rn(0.5, sc);
  }
  bounds[2 * i] = xmin; bounds[2 * i + 1] = xmax;
}

__device__ __forceinline__ int clip8(int v) {
  v >>= RF_PRECISION_BITS;
  return v < 0 ? 0 : (v > 255 ? 255 : v);
}

// Horizontal pass: tmp[b][row][xx][c] (uint8) for row < Q of the padded square.  grid = (row blocks, B), thread = (xx, row).
__global__ void refine_hpass_kernel(const uint8_t* __restrict__ images, const long long* __restrict__ img_off,
                                    const int* __restrict__ img_w, const RefineGeom* __restrict__ geom,
                                    const int* __restrict__ ktab, const int* __restrict__ bounds, int S, int kmax, int qmax,
                                    uint8_t* __restrict__ tmp) {
  const int b = blockIdx.y;
  const RefineGeom g = geom[b];
  if (!g.valid) return;
  const int W = img_w[g.img];
  const uint8_t* img = images + img_off[g.img];
  for (int row = blockIdx.x * blockDim.y + threadIdx.y; row < g.Q; row += gridDim.x * blockDim.y) {
    const int sy = row - g.padt;  // row of the crop (outside [0, ch): grey padding)
    const bool row_in = sy >= 0 && sy < g.ch;
    const uint8_t* src = img + ((long long)(g.iy1 + (row_in ? sy : 0)) * W + g.ix1) * 3;
    for (int xx = threadIdx.x; xx < S; xx += blockDim.x) {
      const int i = b * S + xx;
      const int xmin = bounds[2 * i], cnt = bounds[2 * i + 1];
      const int* k = ktab + (long long)i * kmax;
      int s0 = 1 << (RF_PRECISION_BITS - 1), s1 = s0, s2 = s0;
      for (int x = 0; x < cnt; ++x) {
        const int sx = x + xmin - g.padl;
        int p0 = 128, p1 = 128, p2 = 128;
        if (row_in && sx >= 0 && sx < g.cw) { p0 = src[sx * 3]; p1 = src[sx * 3 + 1]; p2 = src[sx * 3 + 2]; }
        const int kk = k[x];
        s0 += p0 * kk; s1 += p1 * kk; s2 += p2 * kk;
      }
      uint8_t* o = tmp + (((long long)b * qmax + row) * S + xx) * 3;
      o[0] = (uint8_t)clip8(s0); o[1] = (uint8_t)clip8(s1); o[2] = (uint8_t)clip8(s2);
    }
  }
}

// Vertical pass + ToTensor + Normalize(0.5, 0.5): out[b][c][yy][xx] = (v / 255 - 0.5) / 0.5 (fp32 NCHW); u8 (nullable) gets
// the uint8 HWC patch.  The float arithmetic mirrors torchvision: v.float().div(255), then sub(0.5), div(0.5).
__global__ void refine_vpass_kernel(const RefineGeom* __restrict__ geom, const int* __restrict__ ktab, const int* __restrict__ bounds,
                                    int S, int kmax, int qmax, const uint8_t* __restrict__ tmp, const float* __restrict__ fallback,
                                    float* __restrict__ out, uint8_t* __restrict__ u8) {
  const int b = blockIdx.z, yy = blockIdx.y;
  const int xx = blockIdx.x * blockDim.x + threadIdx.x;
  if (xx >= S) return;
  const RefineGeom g = geom[b];
  float* o = out + (long long)b * 3 * S * S + (long long)yy * S + xx;
  if (!g.valid) {  // the reference's except branch (:118-126): the fallback patch, else a zero patch
    const float* f = fallback ? fallback + (long long)b * 3 * S * S + (long long)yy * S + xx : nullptr;
    o[0] = f ? f[0] : 0.f; o[(long long)S * S] = f ? f[(long long)S * S] : 0.f; o[2ll * S * S] = f ? f[2ll * S * S] : 0.f;
    if (u8) { uint8_t* q = u8 + (((long long)b * S + yy) * S + xx) * 3; q[0] = q[1] = q[2] = 0; }
    return;
  }
  const int i = b * S + yy;
  const int ymin = bounds[2 * i], cnt = bounds[2 * i + 1];
  const int* k = ktab + (long long)i * kmax;
  int s0 = 1 << (RF_PRECISION_BITS - 1), s1 = s0, s2 = s0;
  const uint8_t* src = tmp + (((long long)b * qmax + ymin) * S + xx) * 3;
  for (int y = 0; y < cnt; ++y) {
    const int kk = k[y];
    s0 += src[0] * kk; s1 += src[1] * kk; s2 += src[2] * kk;
    src += (long long)S * 3;
  }
  const int v[3] = {clip8(s0), clip8(s1), clip8(s2)};
#pragma unroll
  for (int c = 0; c < 3; ++c) o[(long long)c * S * S] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)v[c], 255.f), 0.5f), 0.5f);
  if (u8) { uint8_t* q = u8 + (((long long)b * S + yy) * S + xx) * 3; q[0] = (uint8_t)v[0]; q[1] = (uint8_t)v[1]; q[2] = (uint8_t)v[2]; }
}
}  // namespace

// Scratch (bytes) for b2c_refine_patches: geometry + coefficient tables + the uint8 intermediate of the horizontal pass.
extern "C" long long b2c_refine_ws_bytes(int B, int S, int qmax) {
  if (B <= 0 || S <= 0 || qmax <= 0) return 0;
  const long long kmax = (long long)((2 * (long long)qmax + S - 1) / S + 1) * 2 + 1;
  long long n = 256 + (long long)B * sizeof(RefineGeom);
  n = (n + 255) / 256 * 256 + (long long)B * S * kmax * 4;
  n = (n + 255) / 256 * 256 + (long long)B * S * 2 * 4;
  n = (n + 255) / 256 * 256 + (long long)B * qmax * S * 3;
  return n + 256;
}

// images: all uint8 HWC RGB images back to back; img_off/img_h/img_w: per image byte offset and size; sample_img[B]: image of
// each sample; pred_box [B][4] normalised (cx,cy,w,h); delta [B][4] or null; mode 0 = dataset letterbox, 1 = refined patch
// (delta applied, clamps, fallback), 2 = inference crop (rounded coordinates); qmax >= the largest image side; ws/ws_bytes:
// b2c_refine_ws_bytes scratch; fallback (nullable, fp32 (B,3,S,S)): patch used for a degenerate crop.  out: fp32 (B,3,S,S); out_u8 (nullable): uint8 (B,S,S,3); box_out (nullable): the cropped boxes.
extern "C" int b2c_refine_patches(const uint8_t* images, const long long* img_off, const int* img_h, const int* img_w,
                                  const int* sample_img, const float* pred_box, const float* delta, int mode,
                                  const float* fallback, void* ws, long long ws_bytes, float* out, uint8_t* out_u8,
                                  float* box_out, int B, int S, int qmax, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  B2C_REQUIRE(mode >= 0 && mode <= 2, "refine_patches: mode=%d", mode);
  B2C_REQUIRE(B > 0 && S > 0 && qmax > 0, "refine_patches: B=%d S=%d qmax=%d", B, S, qmax);
  B2C_REQUIRE(ws && ws_bytes >= b2c_refine_ws_bytes(B, S, qmax), "refine_patches: workspace of %lld bytes < %lld", ws_bytes,
              b2c_refine_ws_bytes(B, S, qmax));
  const int kmax = (int)((2 * (long long)qmax + S - 1) / S + 1) * 2 + 1;
  uint8_t* p = (uint8_t*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
  RefineGeom* geom = (RefineGeom*)p; p += ((size_t)B * sizeof(RefineGeom) + 255) / 256 * 256;
  int* ktab = (int*)p; p += ((size_t)B * S * kmax * 4 + 255) / 256 * 256;
  int* bounds = (int*)p; p += ((size_t)B * S * 2 * 4 + 255) / 256 * 256;
  uint8_t* tmp = p;
  refine_geom_kernel<<<(B + 63) / 64, 64, 0, st>>>(img_h, img_w, sample_img, pred_box, delta, mode, S, geom, box_out, B, kmax);
  refine_coeff_kernel<<<(B * S + 127) / 128, 128, 0, st>>>(geom, S, kmax, ktab, bounds, B);
  const int tx = S < 128 ? S : 128;
  dim3 hb(tx, 256 / tx);
  refine_hpass_kernel<<<dim3((qmax + hb.y - 1) / hb.y > 64 ? 64 : (qmax + hb.y - 1) / hb.y, B), hb, 0, st>>>(
      images, img_off, img_w, geom, ktab, bounds, S, kmax, qmax, tmp);
  refine_vpass_kernel<<<dim3((S + tx - 1) / tx, S, B), tx, 0, st>>>(geom, ktab, bounds, S, kmax, qmax, tmp, fallback, out, out_u8);
  return b2c_check_launch("refine_patches");
}
