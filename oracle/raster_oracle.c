/*
 * raster_oracle.c -- CPU restatement of the reference rasterizer, forward and backward.
 *
 * TEST INFRASTRUCTURE ONLY: used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * reference legs as the checker.  The product (gaussianprediction_b200/) never links or calls it.
 *
 * Parity status: PINNED against outputs of the reference CUDA build (oracle/_ref, built from
 * /root/reference by oracle/build_ref.py) run on a B200; the vectors are tests/golden/raster_*.npz,
 * produced by tests/golden/make_golden.py.  The reference has no CPU rasterizer and no tests.
 *
 * Each function cites the reference code it follows, paths relative to
 * submodules/diff-gaussian-rasterization-w-depth/cuda_rasterizer/.
 * Plain single-threaded C, FP32 arithmetic, built with -ffp-contract=off: it is an independent
 * evaluation of the same formulas, so it agrees with the CUDA results to rounding, not bit for bit.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define BX 16
#define BY 16

typedef struct { float x, y, z; } v3;

static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                               -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

/* auxiliary.h:58-99 */
static v3 xform4x3(v3 p, const float* m) {
  v3 r = {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
          m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
  return r;
}
static void xform4x4(v3 p, const float* m, float* o) {
  o[0] = m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12];
  o[1] = m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13];
  o[2] = m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14];
  o[3] = m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15];
}
/* auxiliary.h:41-44 (double arithmetic: unsuffixed literals) */
static float ndc2pix(float v, int S) { return (float)(((v + 1.0) * S - 1.0) * 0.5); }

static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

/* auxiliary.h:46-56 */
static void get_rect(float px, float py, int r, int gx, int gy, int* x0, int* y0, int* x1, int* y1) {
  *x0 = imin(gx, imax(0, (int)((px - r) / BX)));
  *y0 = imin(gy, imax(0, (int)((py - r) / BY)));
  *x1 = imin(gx, imax(0, (int)((px + r + BX - 1) / BX)));
  *y1 = imin(gy, imax(0, (int)((py + r + BY - 1) / BY)));
}

/* 3x3 helpers, column-major m[col*3+row], product order as the reference's matrix library */
static void m3mul(const float* a, const float* b, float* r) {
  for (int j = 0; j < 3; j++)
    for (int i = 0; i < 3; i++) r[j * 3 + i] = a[0 * 3 + i] * b[j * 3 + 0] + a[1 * 3 + i] * b[j * 3 + 1] + a[2 * 3 + i] * b[j * 3 + 2];
}
static void m3t(const float* a, float* r) {
  for (int j = 0; j < 3; j++)
    for (int i = 0; i < 3; i++) r[j * 3 + i] = a[i * 3 + j];
}

/* forward.cu:118-152 (quaternion NOT normalised in this fork, :127) */
static void cov3d(const float* s, float mod, const float* q, float* out) {
  float S[9] = {mod * s[0], 0, 0, 0, mod * s[1], 0, 0, 0, mod * s[2]};
  float r = q[0], x = q[1], y = q[2], z = q[3];
  float R[9] = {1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)};
  float M[9], Mt[9], Sg[9];
  m3mul(S, R, M);
  m3t(M, Mt);
  m3mul(Mt, M, Sg);
  out[0] = Sg[0]; out[1] = Sg[1]; out[2] = Sg[2]; out[3] = Sg[4]; out[4] = Sg[5]; out[5] = Sg[8];
}

/* forward.cu:74-113; also returns T (for the backward) and the clamped t */
static void cov2d(v3 mean, float fx, float fy, float tanx, float tany, const float* c3, const float* view,
                  float* cov /*3*/, float* Tout /*9 or NULL*/, v3* tout, float* xmul, float* ymul) {
  v3 t = xform4x3(mean, view);
  float limx = 1.3f * tanx, limy = 1.3f * tany;
  float txtz = t.x / t.z, tytz = t.y / t.z;
  t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
  t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
  if (xmul) *xmul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
  if (ymul) *ymul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
  float J[9] = {fx / t.z, 0.f, -(fx * t.x) / (t.z * t.z), 0.f, fy / t.z, -(fy * t.y) / (t.z * t.z), 0, 0, 0};
  float W[9] = {view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]};
  float T[9], Tt[9], V[9] = {c3[0], c3[1], c3[2], c3[1], c3[3], c3[4], c3[2], c3[4], c3[5]}, Vt[9], A[9], C[9];
  m3mul(W, J, T);
  m3t(T, Tt);
  m3t(V, Vt);
  m3mul(Tt, Vt, A);
  m3mul(A, T, C);
  cov[0] = C[0] + 0.3f;
  cov[1] = C[1];
  cov[2] = C[4] + 0.3f;
  if (Tout) memcpy(Tout, T, sizeof(T));
  if (tout) *tout = t;
}

/* forward.cu:20-71 */
static void sh2rgb(int deg, v3 pos, const float* campos, const float* sh, float* rgb, uint8_t* clamped) {
  v3 d = {pos.x - campos[0], pos.y - campos[1], pos.z - campos[2]};
  float len = sqrtf(d.x * d.x + d.y * d.y + d.z * d.z);
  float x = d.x / len, y = d.y / len, z = d.z / len;
  for (int c = 0; c < 3; c++) {
#define S(k) sh[(k) * 3 + c]
    float r = SH_C0 * S(0);
    if (deg > 0) {
      r = r - SH_C1 * y * S(1) + SH_C1 * z * S(2) - SH_C1 * x * S(3);
      if (deg > 1) {
        float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        r = r + SH_C2[0] * xy * S(4) + SH_C2[1] * yz * S(5) + SH_C2[2] * (2.0f * zz - xx - yy) * S(6) +
            SH_C2[3] * xz * S(7) + SH_C2[4] * (xx - yy) * S(8);
        if (deg > 2) {
          r = r + SH_C3[0] * y * (3.0f * xx - yy) * S(9) + SH_C3[1] * xy * z * S(10) +
              SH_C3[2] * y * (4.0f * zz - xx - yy) * S(11) + SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S(12) +
              SH_C3[4] * x * (4.0f * zz - xx - yy) * S(13) + SH_C3[5] * z * (xx - yy) * S(14) +
              SH_C3[6] * x * (xx - 3.0f * yy) * S(15);
        }
      }
    }
#undef S
    r += 0.5f;
    clamped[c] = r < 0;
    rgb[c] = r > 0.f ? r : 0.f;
  }
}

typedef struct { uint64_t key; uint32_t val; } kv_t;
static int kv_cmp(const void* a, const void* b) {
  const kv_t* p = (const kv_t*)a; const kv_t* q = (const kv_t*)b;
  if (p->key != q->key) return p->key < q->key ? -1 : 1;
  return p->val < q->val ? -1 : (p->val > q->val);   /* stable radix sort of index-ordered emission == tie by index */
}

/*
 * Forward.  Follows Rasterizer::forward (rasterizer_impl.cu:198-342): preprocessCUDA (forward.cu:155-256),
 * inclusive scan, duplicateWithKeys (rasterizer_impl.cu:70-111), stable sort by (tile<<32 | depth bits),
 * identifyTileRanges (:116-138), renderCUDA (forward.cu:261-398).
 * Per-Gaussian outputs are left at 0 for culled Gaussians.  point_list/keys are malloc'ed here and handed
 * back through *point_list_out / *keys_out (free with oracle_free).  Returns num_rendered.
 */
long long oracle_raster_forward(int P, int D, int M, const float* bg, int W, int H, const float* means3D,
                                const float* shs, const float* colors_precomp, const float* opacities,
                                const float* scales, float scale_modifier, const float* rotations,
                                const float* cov3D_precomp, const float* view, const float* proj, const float* campos,
                                float tanx, float tany,
                                /* per-Gaussian state */
                                int* radii, float* depths, float* means2D, float* cov3Ds, float* conic_opacity,
                                float* rgb, uint32_t* tiles_touched, uint8_t* clamped,
                                /* binning */
                                uint32_t** point_list_out, uint64_t** keys_out, uint32_t* ranges /*[tiles*2]*/,
                                /* image */
                                float* out_color, float* out_depth, int* out_tidx, float* final_T, uint32_t* n_contrib) {
  const float fy = H / (2.0f * tany), fx = W / (2.0f * tanx);
  const int gx = (W + BX - 1) / BX, gy = (H + BY - 1) / BY;
  long long R = 0;
  for (int i = 0; i < P; i++) {
    radii[i] = 0; tiles_touched[i] = 0;
    v3 p = {means3D[3 * i], means3D[3 * i + 1], means3D[3 * i + 2]};
    v3 pv = xform4x3(p, view);
    if (pv.z <= 0.2f) continue;                                   /* auxiliary.h:154 */
    float ph[4]; xform4x4(p, proj, ph);
    float pw = 1.0f / (ph[3] + 0.0000001f);
    float ppx = ph[0] * pw, ppy = ph[1] * pw;
    const float* c3;
    if (cov3D_precomp) c3 = cov3D_precomp + 6 * (size_t)i;
    else { cov3d(scales + 3 * (size_t)i, scale_modifier, rotations + 4 * (size_t)i, cov3Ds + 6 * (size_t)i); c3 = cov3Ds + 6 * (size_t)i; }
    float cov[3];
    cov2d(p, fx, fy, tanx, tany, c3, view, cov, NULL, NULL, NULL, NULL);
    float det = cov[0] * cov[2] - cov[1] * cov[1];
    if (det == 0.0f) continue;
    float det_inv = 1.f / det;
    float conic[3] = {cov[2] * det_inv, -cov[1] * det_inv, cov[0] * det_inv};
    float mid = 0.5f * (cov[0] + cov[2]);
    float l1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
    float l2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
    float my_radius = ceilf(3.f * sqrtf(fmaxf(l1, l2)));
    float pix = ndc2pix(ppx, W), piy = ndc2pix(ppy, H);
    int x0, y0, x1, y1;
    get_rect(pix, piy, (int)my_radius, gx, gy, &x0, &y0, &x1, &y1);
    if ((x1 - x0) * (y1 - y0) == 0) continue;
    if (!colors_precomp) sh2rgb(D, p, campos, shs + (size_t)i * M * 3, rgb + 3 * (size_t)i, clamped + 3 * (size_t)i);
    depths[i] = pv.z;
    radii[i] = (int)my_radius;
    means2D[2 * i] = pix; means2D[2 * i + 1] = piy;
    conic_opacity[4 * i] = conic[0]; conic_opacity[4 * i + 1] = conic[1]; conic_opacity[4 * i + 2] = conic[2];
    conic_opacity[4 * i + 3] = opacities[i];
    tiles_touched[i] = (uint32_t)((y1 - y0) * (x1 - x0));
    R += tiles_touched[i];
  }
  kv_t* kv = (kv_t*)malloc(sizeof(kv_t) * (size_t)(R ? R : 1));
  size_t off = 0;
  for (int i = 0; i < P; i++) {
    if (radii[i] <= 0) continue;
    int x0, y0, x1, y1;
    get_rect(means2D[2 * i], means2D[2 * i + 1], radii[i], gx, gy, &x0, &y0, &x1, &y1);
    uint32_t dbits; memcpy(&dbits, &depths[i], 4);
    for (int y = y0; y < y1; y++)
      for (int x = x0; x < x1; x++) {
        kv[off].key = ((uint64_t)(y * gx + x) << 32) | dbits;
        kv[off].val = (uint32_t)i;
        off++;
      }
  }
  qsort(kv, (size_t)R, sizeof(kv_t), kv_cmp);
  uint32_t* plist = (uint32_t*)malloc(4 * (size_t)(R ? R : 1));
  uint64_t* keys = (uint64_t*)malloc(8 * (size_t)(R ? R : 1));
  for (long long i = 0; i < R; i++) { plist[i] = kv[i].val; keys[i] = kv[i].key; }
  free(kv);
  memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)gx * gy);
  for (long long i = 0; i < R; i++) {
    uint32_t cur = (uint32_t)(keys[i] >> 32);
    if (i == 0) ranges[2 * cur] = 0;
    else { uint32_t prev = (uint32_t)(keys[i - 1] >> 32); if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)i; ranges[2 * cur] = (uint32_t)i; } }
    if (i == R - 1) ranges[2 * cur + 1] = (uint32_t)R;
  }
  const float* feat = colors_precomp ? colors_precomp : rgb;
  for (int py = 0; py < H; py++)
    for (int px = 0; px < W; px++) {
      int tile = (py / BY) * gx + (px / BX);
      uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
      float T = 1.0f, C[3] = {0, 0, 0}, Dm = 15.0f;               /* forward.cu:310 sentinel */
      int tid = -1; uint32_t contributor = 0, last = 0;
      float pfx = (float)px, pfy = (float)py;
      for (uint32_t k = r0; k < r1; k++) {
        contributor++;
        uint32_t g = plist[k];
        float dx = means2D[2 * g] - pfx, dy = means2D[2 * g + 1] - pfy;
        const float* co = conic_opacity + 4 * (size_t)g;
        float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
        if (power > 0.0f) continue;
        float alpha = fminf(0.99f, co[3] * expf(power));
        if (alpha < 1.0f / 255.0f) continue;
        float test_T = T * (1 - alpha);
        if (test_T < 0.0001f) break;                               /* done = true */
        for (int ch = 0; ch < 3; ch++) C[ch] += feat[3 * (size_t)g + ch] * alpha * T;
        if (T > 0.5f && test_T < 0.5f) { Dm = depths[g]; tid = (int)g; }   /* median depth, forward.cu:371-376 */
        T = test_T;
        last = contributor;
      }
      size_t pid = (size_t)W * py + px;
      final_T[pid] = T; n_contrib[pid] = last;
      for (int ch = 0; ch < 3; ch++) out_color[(size_t)ch * H * W + pid] = C[ch] + T * bg[ch];
      out_depth[pid] = Dm; out_tidx[pid] = tid;
    }
  *point_list_out = plist; *keys_out = keys;
  return R;
}

void oracle_free(void* p) { free(p); }

/* auxiliary.h:101-113 */
static v3 dnormvdv(v3 v, v3 dv) {
  float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
  float inv = 1.0f / sqrtf(sum2 * sum2 * sum2);
  v3 r;
  r.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * inv;
  r.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * inv;
  r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * inv;
  return r;
}

/*
 * Backward.  Follows Rasterizer::backward (rasterizer_impl.cu:346-439): renderCUDA (backward.cu:399-557)
 * with the atomics replaced by plain accumulation, computeCov2DCUDA (:144-274), preprocessCUDA (:346-396)
 * incl. SH backward (:20-139) and computeCov3D backward (:278-341).
 * All outputs must be zero-initialised by the caller (as rasterize_points.cu:153-161 does).
 */
void oracle_raster_backward(int P, int D, int M, const float* bg, int W, int H, const float* means3D, const float* shs,
                            const float* colors_precomp, const float* scales, float scale_modifier,
                            const float* rotations, const float* cov3D_precomp, const float* view, const float* proj,
                            const float* campos, float tanx, float tany, const int* radii, const float* means2D,
                            const float* cov3Ds, const float* conic_opacity, const float* rgb, const uint8_t* clamped,
                            const uint32_t* point_list, const uint32_t* ranges, const float* final_T,
                            const uint32_t* n_contrib, const float* dL_dpix,
                            float* dL_dmean2D /*[P,3]*/, float* dL_dconic /*[P,4]*/, float* dL_dopacity /*[P]*/,
                            float* dL_dcolor /*[P,3]*/, float* dL_dmean3D /*[P,3]*/, float* dL_dcov3D /*[P,6]*/,
                            float* dL_dsh /*[P,M,3]*/, float* dL_dscale /*[P,3]*/, float* dL_drot /*[P,4]*/) {
  const float fy = H / (2.0f * tany), fx = W / (2.0f * tanx);
  const int gx = (W + BX - 1) / BX;
  const float* colors = colors_precomp ? colors_precomp : rgb;
  const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
  for (int py = 0; py < H; py++)
    for (int px = 0; px < W; px++) {
      int tile = (py / BY) * gx + (px / BX);
      uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
      size_t pid = (size_t)W * py + px;
      const float T_final = final_T[pid];
      float T = T_final;
      uint32_t contributor = r1 - r0;
      const uint32_t last = n_contrib[pid];
      float accum[3] = {0, 0, 0}, dpix[3], last_alpha = 0, last_color[3] = {0, 0, 0};
      for (int c = 0; c < 3; c++) dpix[c] = dL_dpix[(size_t)c * H * W + pid];
      float pfx = (float)px, pfy = (float)py;
      for (uint32_t k = r1; k-- > r0;) {                       /* back to front */
        contributor--;
        if (contributor >= last) continue;
        uint32_t g = point_list[k];
        float dx = means2D[2 * g] - pfx, dy = means2D[2 * g + 1] - pfy;
        const float* co = conic_opacity + 4 * (size_t)g;
        float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
        if (power > 0.0f) continue;
        float G = expf(power);
        float alpha = fminf(0.99f, co[3] * G);
        if (alpha < 1.0f / 255.0f) continue;
        T = T / (1.f - alpha);
        float dchannel_dcolor = alpha * T;
        float dL_dalpha = 0.0f;
        for (int ch = 0; ch < 3; ch++) {
          float c = colors[3 * (size_t)g + ch];
          accum[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum[ch];
          last_color[ch] = c;
          dL_dalpha += (c - accum[ch]) * dpix[ch];
          dL_dcolor[3 * (size_t)g + ch] += dchannel_dcolor * dpix[ch];
        }
        dL_dalpha *= T;
        last_alpha = alpha;
        float bgdot = 0;
        for (int c = 0; c < 3; c++) bgdot += bg[c] * dpix[c];
        dL_dalpha += (-T_final / (1.f - alpha)) * bgdot;
        float dL_dG = co[3] * dL_dalpha;
        float gdx = G * dx, gdy = G * dy;
        float dG_ddelx = -gdx * co[0] - gdy * co[1];
        float dG_ddely = -gdy * co[2] - gdx * co[1];
        dL_dmean2D[3 * (size_t)g] += dL_dG * dG_ddelx * ddelx_dx;
        dL_dmean2D[3 * (size_t)g + 1] += dL_dG * dG_ddely * ddely_dy;
        dL_dconic[4 * (size_t)g] += -0.5f * gdx * dx * dL_dG;
        dL_dconic[4 * (size_t)g + 1] += -0.5f * gdx * dy * dL_dG;
        dL_dconic[4 * (size_t)g + 3] += -0.5f * gdy * dy * dL_dG;
        dL_dopacity[g] += G * dL_dalpha;
      }
    }
  for (int i = 0; i < P; i++) {
    if (!(radii[i] > 0)) continue;
    v3 mean = {means3D[3 * (size_t)i], means3D[3 * (size_t)i + 1], means3D[3 * (size_t)i + 2]};
    const float* c3 = cov3D_precomp ? cov3D_precomp + 6 * (size_t)i : cov3Ds + 6 * (size_t)i;
    /* computeCov2DCUDA */
    float cov[3], T[9], xmul, ymul; v3 t;
    cov2d(mean, fx, fy, tanx, tany, c3, view, cov, T, &t, &xmul, &ymul);
    float a = cov[0], b = cov[1], c = cov[2];
    float dcx = dL_dconic[4 * (size_t)i], dcy = dL_dconic[4 * (size_t)i + 1], dcw = dL_dconic[4 * (size_t)i + 3];
    float denom = a * c - b * b, dL_da = 0, dL_db = 0, dL_dc = 0;
    float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float* dcv = dL_dcov3D + 6 * (size_t)i;
#define TT(col, row) T[(col) * 3 + (row)]
    if (denom2inv != 0) {
      dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcw);
      dL_dc = denom2inv * (-a * a * dcw + 2 * a * b * dcy + (denom - a * c) * dcx);
      dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcw);
      dcv[0] = (TT(0, 0) * TT(0, 0) * dL_da + TT(0, 0) * TT(1, 0) * dL_db + TT(1, 0) * TT(1, 0) * dL_dc);
      dcv[3] = (TT(0, 1) * TT(0, 1) * dL_da + TT(0, 1) * TT(1, 1) * dL_db + TT(1, 1) * TT(1, 1) * dL_dc);
      dcv[5] = (TT(0, 2) * TT(0, 2) * dL_da + TT(0, 2) * TT(1, 2) * dL_db + TT(1, 2) * TT(1, 2) * dL_dc);
      dcv[1] = 2 * TT(0, 0) * TT(0, 1) * dL_da + (TT(0, 0) * TT(1, 1) + TT(0, 1) * TT(1, 0)) * dL_db + 2 * TT(1, 0) * TT(1, 1) * dL_dc;
      dcv[2] = 2 * TT(0, 0) * TT(0, 2) * dL_da + (TT(0, 0) * TT(1, 2) + TT(0, 2) * TT(1, 0)) * dL_db + 2 * TT(1, 0) * TT(1, 2) * dL_dc;
      dcv[4] = 2 * TT(0, 2) * TT(0, 1) * dL_da + (TT(0, 1) * TT(1, 2) + TT(0, 2) * TT(1, 1)) * dL_db + 2 * TT(1, 1) * TT(1, 2) * dL_dc;
    }
    float V[9] = {c3[0], c3[1], c3[2], c3[1], c3[3], c3[4], c3[2], c3[4], c3[5]};
#define VV(col, row) V[(col) * 3 + (row)]
    float dT00 = 2 * (TT(0, 0) * VV(0, 0) + TT(0, 1) * VV(0, 1) + TT(0, 2) * VV(0, 2)) * dL_da + (TT(1, 0) * VV(0, 0) + TT(1, 1) * VV(0, 1) + TT(1, 2) * VV(0, 2)) * dL_db;
    float dT01 = 2 * (TT(0, 0) * VV(1, 0) + TT(0, 1) * VV(1, 1) + TT(0, 2) * VV(1, 2)) * dL_da + (TT(1, 0) * VV(1, 0) + TT(1, 1) * VV(1, 1) + TT(1, 2) * VV(1, 2)) * dL_db;
    float dT02 = 2 * (TT(0, 0) * VV(2, 0) + TT(0, 1) * VV(2, 1) + TT(0, 2) * VV(2, 2)) * dL_da + (TT(1, 0) * VV(2, 0) + TT(1, 1) * VV(2, 1) + TT(1, 2) * VV(2, 2)) * dL_db;
    float dT10 = 2 * (TT(1, 0) * VV(0, 0) + TT(1, 1) * VV(0, 1) + TT(1, 2) * VV(0, 2)) * dL_dc + (TT(0, 0) * VV(0, 0) + TT(0, 1) * VV(0, 1) + TT(0, 2) * VV(0, 2)) * dL_db;
    float dT11 = 2 * (TT(1, 0) * VV(1, 0) + TT(1, 1) * VV(1, 1) + TT(1, 2) * VV(1, 2)) * dL_dc + (TT(0, 0) * VV(1, 0) + TT(0, 1) * VV(1, 1) + TT(0, 2) * VV(1, 2)) * dL_db;
    float dT12 = 2 * (TT(1, 0) * VV(2, 0) + TT(1, 1) * VV(2, 1) + TT(1, 2) * VV(2, 2)) * dL_dc + (TT(0, 0) * VV(2, 0) + TT(0, 1) * VV(2, 1) + TT(0, 2) * VV(2, 2)) * dL_db;
#undef VV
#undef TT
    float Wm[9] = {view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]};
#define WW(col, row) Wm[(col) * 3 + (row)]
    float dJ00 = WW(0, 0) * dT00 + WW(0, 1) * dT01 + WW(0, 2) * dT02;
    float dJ02 = WW(2, 0) * dT00 + WW(2, 1) * dT01 + WW(2, 2) * dT02;
    float dJ11 = WW(1, 0) * dT10 + WW(1, 1) * dT11 + WW(1, 2) * dT12;
    float dJ12 = WW(2, 0) * dT10 + WW(2, 1) * dT11 + WW(2, 2) * dT12;
#undef WW
    float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
    float dtx = xmul * -fx * tz2 * dJ02;
    float dty = ymul * -fy * tz2 * dJ12;
    float dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2 * fx * t.x) * tz3 * dJ02 + (2 * fy * t.y) * tz3 * dJ12;
    v3 dm = {view[0] * dtx + view[1] * dty + view[2] * dtz, view[4] * dtx + view[5] * dty + view[6] * dtz,
             view[8] * dtx + view[9] * dty + view[10] * dtz};
    /* preprocessCUDA backward: projection part */
    float mh[4]; xform4x4(mean, proj, mh);
    float mw = 1.0f / (mh[3] + 0.0000001f);
    float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * mw * mw;
    float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * mw * mw;
    float g2x = dL_dmean2D[3 * (size_t)i], g2y = dL_dmean2D[3 * (size_t)i + 1];
    dm.x += (proj[0] * mw - proj[3] * mul1) * g2x + (proj[1] * mw - proj[3] * mul2) * g2y;
    dm.y += (proj[4] * mw - proj[7] * mul1) * g2x + (proj[5] * mw - proj[7] * mul2) * g2y;
    dm.z += (proj[8] * mw - proj[11] * mul1) * g2x + (proj[9] * mw - proj[11] * mul2) * g2y;
    /* SH backward */
    if (shs) {
      const float* sh = shs + (size_t)i * M * 3;
      float* dsh = dL_dsh + (size_t)i * M * 3;
      v3 dorig = {mean.x - campos[0], mean.y - campos[1], mean.z - campos[2]};
      float len = sqrtf(dorig.x * dorig.x + dorig.y * dorig.y + dorig.z * dorig.z);
      float x = dorig.x / len, y = dorig.y / len, z = dorig.z / len;
      float dRGB[3], ddx = 0, ddy = 0, ddz = 0;
      for (int ch = 0; ch < 3; ch++) dRGB[ch] = dL_dcolor[3 * (size_t)i + ch] * (clamped[3 * (size_t)i + ch] ? 0.f : 1.f);
      for (int ch = 0; ch < 3; ch++) {
#define S(k) sh[(k) * 3 + ch]
#define DS(k) dsh[(k) * 3 + ch]
        float g = dRGB[ch], rx = 0, ry = 0, rz = 0;
        DS(0) = SH_C0 * g;
        if (D > 0) {
          DS(1) = -SH_C1 * y * g; DS(2) = SH_C1 * z * g; DS(3) = -SH_C1 * x * g;
          rx = -SH_C1 * S(3); ry = -SH_C1 * S(1); rz = SH_C1 * S(2);
          if (D > 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            DS(4) = SH_C2[0] * xy * g; DS(5) = SH_C2[1] * yz * g; DS(6) = SH_C2[2] * (2.f * zz - xx - yy) * g;
            DS(7) = SH_C2[3] * xz * g; DS(8) = SH_C2[4] * (xx - yy) * g;
            rx += SH_C2[0] * y * S(4) + SH_C2[2] * 2.f * -x * S(6) + SH_C2[3] * z * S(7) + SH_C2[4] * 2.f * x * S(8);
            ry += SH_C2[0] * x * S(4) + SH_C2[1] * z * S(5) + SH_C2[2] * 2.f * -y * S(6) + SH_C2[4] * 2.f * -y * S(8);
            rz += SH_C2[1] * y * S(5) + SH_C2[2] * 2.f * 2.f * z * S(6) + SH_C2[3] * x * S(7);
            if (D > 2) {
              DS(9) = SH_C3[0] * y * (3.f * xx - yy) * g; DS(10) = SH_C3[1] * xy * z * g;
              DS(11) = SH_C3[2] * y * (4.f * zz - xx - yy) * g; DS(12) = SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy) * g;
              DS(13) = SH_C3[4] * x * (4.f * zz - xx - yy) * g; DS(14) = SH_C3[5] * z * (xx - yy) * g;
              DS(15) = SH_C3[6] * x * (xx - 3.f * yy) * g;
              rx += SH_C3[0] * S(9) * 3.f * 2.f * xy + SH_C3[1] * S(10) * yz + SH_C3[2] * S(11) * -2.f * xy +
                    SH_C3[3] * S(12) * -3.f * 2.f * xz + SH_C3[4] * S(13) * (-3.f * xx + 4.f * zz - yy) +
                    SH_C3[5] * S(14) * 2.f * xz + SH_C3[6] * S(15) * 3.f * (xx - yy);
              ry += SH_C3[0] * S(9) * 3.f * (xx - yy) + SH_C3[1] * S(10) * xz + SH_C3[2] * S(11) * (-3.f * yy + 4.f * zz - xx) +
                    SH_C3[3] * S(12) * -3.f * 2.f * yz + SH_C3[4] * S(13) * -2.f * xy + SH_C3[5] * S(14) * -2.f * yz +
                    SH_C3[6] * S(15) * -3.f * 2.f * xy;
              rz += SH_C3[1] * S(10) * xy + SH_C3[2] * S(11) * 4.f * 2.f * yz + SH_C3[3] * S(12) * 3.f * (2.f * zz - xx - yy) +
                    SH_C3[4] * S(13) * 4.f * 2.f * xz + SH_C3[5] * S(14) * (xx - yy);
            }
          }
        }
#undef S
#undef DS
        ddx += rx * g; ddy += ry * g; ddz += rz * g;
      }
      v3 dd = {ddx, ddy, ddz};
      v3 dmm = dnormvdv(dorig, dd);
      dm.x += dmm.x; dm.y += dmm.y; dm.z += dmm.z;
    }
    dL_dmean3D[3 * (size_t)i] = dm.x; dL_dmean3D[3 * (size_t)i + 1] = dm.y; dL_dmean3D[3 * (size_t)i + 2] = dm.z;
    /* computeCov3D backward (raw quaternion) */
    if (scales) {
      const float* q = rotations + 4 * (size_t)i;
      float r = q[0], x = q[1], y = q[2], z = q[3];
      float Rm[9] = {1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                     2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                     2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)};
      float s[3] = {scale_modifier * scales[3 * (size_t)i], scale_modifier * scales[3 * (size_t)i + 1], scale_modifier * scales[3 * (size_t)i + 2]};
      float S[9] = {s[0], 0, 0, 0, s[1], 0, 0, 0, s[2]}, Mm[9], M2[9], dSig[9], dM[9], Rt[9], dMt[9];
      m3mul(S, Rm, Mm);
      dSig[0] = dcv[0]; dSig[1] = 0.5f * dcv[1]; dSig[2] = 0.5f * dcv[2];
      dSig[3] = 0.5f * dcv[1]; dSig[4] = dcv[3]; dSig[5] = 0.5f * dcv[4];
      dSig[6] = 0.5f * dcv[2]; dSig[7] = 0.5f * dcv[4]; dSig[8] = dcv[5];
      for (int k = 0; k < 9; k++) M2[k] = Mm[k] * 2.0f;
      m3mul(M2, dSig, dM);
      m3t(Rm, Rt);
      m3t(dM, dMt);
#define A(col, row) dMt[(col) * 3 + (row)]
      for (int k = 0; k < 3; k++) dL_dscale[3 * (size_t)i + k] = Rt[k * 3 + 0] * A(k, 0) + Rt[k * 3 + 1] * A(k, 1) + Rt[k * 3 + 2] * A(k, 2);
      for (int k = 0; k < 3; k++) { A(0, k) *= s[0]; A(1, k) *= s[1]; A(2, k) *= s[2]; }
      float* dq = dL_drot + 4 * (size_t)i;
      dq[0] = 2 * z * (A(0, 1) - A(1, 0)) + 2 * y * (A(2, 0) - A(0, 2)) + 2 * x * (A(1, 2) - A(2, 1));
      dq[1] = 2 * y * (A(1, 0) + A(0, 1)) + 2 * z * (A(2, 0) + A(0, 2)) + 2 * r * (A(1, 2) - A(2, 1)) - 4 * x * (A(2, 2) + A(1, 1));
      dq[2] = 2 * x * (A(1, 0) + A(0, 1)) + 2 * r * (A(2, 0) - A(0, 2)) + 2 * z * (A(1, 2) + A(2, 1)) - 4 * y * (A(2, 2) + A(0, 0));
      dq[3] = 2 * r * (A(0, 1) - A(1, 0)) + 2 * x * (A(2, 0) + A(0, 2)) + 2 * y * (A(1, 2) + A(2, 1)) - 4 * z * (A(1, 1) + A(0, 0));
#undef A
    }
  }
}
