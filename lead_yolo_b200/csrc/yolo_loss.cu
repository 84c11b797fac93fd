// YOLOv5 detection loss with the EIoU box term — reference utils/loss.py:121-268 (ComputeLoss.__call__,
// build_targets) and utils/metrics.py:293-354 (bbox_iou(..., EIoU=True)) — as a handful of small fp32 kernels with
// static shapes, so that a whole training step can be captured in a CUDA graph (the reference filters its target list
// with boolean masks, i.e. data-dependent shapes and a host synchronisation per level).
//
// Data: raw head maps p_i (bs, na, ny_i, nx_i, no) fp32 with arbitrary element strides (the Detect training branch
// returns permuted views of its NHWC convolution output); targets [nt][6] = (image, class, x, y, w, h) normalised,
// rows with image < 0 are padding.  Candidate "entries" are indexed id = (o*na + a)*nt + t per level (o = the five
// neighbour offsets of build_targets, a = anchor, t = target) — the order of the reference's filtered list, so its
// "last write wins" scatter into tobj (loss.py:171) is the entry with the largest id on a cell.
//
// Objectness over all cells is split into a dense part Σ softplus(x) (target 0 everywhere) and a correction on the
// winning entry of each target cell, so no dense target map is ever built.
#include <math.h>

#include "ly_common.cuh"

namespace ly {

constexpr int kMaxLevels = 3;
constexpr int kRec = 8;  // floats per entry record: cell (int bits) | tx ty tw th | aw ah | cls (int bits); cell < 0 = unused

struct LossGeom {
  int nl, na, nc, no, bs, nt;
  int ny[kMaxLevels], nx[kMaxLevels];
  const float* p[kMaxLevels];
  float* dp[kMaxLevels];
  long sb[kMaxLevels], sa[kMaxLevels], sy[kMaxLevels], sx[kMaxLevels], so[kMaxLevels];
  long cell_off[kMaxLevels];  // first cell of the level in the winner map
  long ent_off[kMaxLevels];   // first entry of the level (entries per level = 5*na*nt)
  float balance[kMaxLevels];
  float anchor_t, box, obj, cls, cls_pw, obj_pw, cp, cn;
};

// ---- forward-mode derivatives with respect to the four decoded box parameters (x, y, w, h) --------------------
struct D4 {
  float v, d[4];
};
__device__ __forceinline__ D4 dconst(float v) { return D4{v, {0.f, 0.f, 0.f, 0.f}}; }
__device__ __forceinline__ D4 dvar(float v, int i) {
  D4 r = dconst(v);
  r.d[i] = 1.f;
  return r;
}
__device__ __forceinline__ D4 operator+(const D4& a, const D4& b) {
  D4 r;
  r.v = a.v + b.v;
#pragma unroll
  for (int i = 0; i < 4; ++i) r.d[i] = a.d[i] + b.d[i];
  return r;
}
__device__ __forceinline__ D4 operator-(const D4& a, const D4& b) {
  D4 r;
  r.v = a.v - b.v;
#pragma unroll
  for (int i = 0; i < 4; ++i) r.d[i] = a.d[i] - b.d[i];
  return r;
}
__device__ __forceinline__ D4 operator*(const D4& a, const D4& b) {
  D4 r;
  r.v = a.v * b.v;
#pragma unroll
  for (int i = 0; i < 4; ++i) r.d[i] = a.d[i] * b.v + a.v * b.d[i];
  return r;
}
__device__ __forceinline__ D4 operator/(const D4& a, const D4& b) {
  D4 r;
  const float inv = 1.f / b.v;
  r.v = a.v * inv;
#pragma unroll
  for (int i = 0; i < 4; ++i) r.d[i] = (a.d[i] - r.v * b.d[i]) * inv;
  return r;
}
__device__ __forceinline__ D4 operator*(const D4& a, float s) {
  D4 r;
  r.v = a.v * s;
#pragma unroll
  for (int i = 0; i < 4; ++i) r.d[i] = a.d[i] * s;
  return r;
}
__device__ __forceinline__ D4 operator+(const D4& a, float s) {
  D4 r = a;
  r.v += s;
  return r;
}
// torch.minimum / maximum / clamp(0): the gradient follows the selected operand (ties: the first, as ATen's
// minimum/maximum backward splits ties evenly only for exact equality, which fp32 boxes do not produce)
__device__ __forceinline__ D4 dmin(const D4& a, const D4& b) { return a.v <= b.v ? a : b; }
__device__ __forceinline__ D4 dmax(const D4& a, const D4& b) { return a.v >= b.v ? a : b; }
__device__ __forceinline__ D4 dclamp0(const D4& a) { return a.v > 0.f ? a : dconst(0.f); }

// bbox_iou(box1 = prediction xywh, box2 = target xywh, EIoU=True), metrics.py:309-347 with alpha = 1, eps = 1e-7
__device__ __forceinline__ D4 eiou_xywh(const D4& x1, const D4& y1, const D4& w1, const D4& h1, float x2, float y2, float w2,
                                        float h2) {
  const float eps = 1e-7f;
  const D4 b1x1 = x1 - w1 * 0.5f, b1x2 = x1 + w1 * 0.5f, b1y1 = y1 - h1 * 0.5f, b1y2 = y1 + h1 * 0.5f;
  const D4 b2x1 = dconst(x2 - w2 / 2), b2x2 = dconst(x2 + w2 / 2), b2y1 = dconst(y2 - h2 / 2), b2y2 = dconst(y2 + h2 / 2);
  const D4 inter = dclamp0(dmin(b1x2, b2x2) - dmax(b1x1, b2x1)) * dclamp0(dmin(b1y2, b2y2) - dmax(b1y1, b2y1));
  const D4 uni = w1 * h1 + (w2 * h2 + eps) - inter;
  const D4 iou = inter / (uni + eps);
  const D4 cw = dmax(b1x2, b2x2) - dmin(b1x1, b2x1);
  const D4 ch = dmax(b1y2, b2y2) - dmin(b1y1, b2y1);
  const D4 cw2 = cw * cw + eps, ch2 = ch * ch + eps;
  const D4 c2 = cw * cw + ch * ch + eps;
  const D4 dx = (b2x1 + b2x2) - b1x1 - b1x2, dy = (b2y1 + b2y2) - b1y1 - b1y2;
  const D4 rho2 = (dx * dx + dy * dy) * 0.25f;
  const D4 dw = (b2x2 - b2x1) - (b1x2 - b1x1), dh = (b2y2 - b2y1) - (b1y2 - b1y1);
  return iou - (rho2 / c2 + (dw * dw) / cw2 + (dh * dh) / ch2);
}

__device__ __forceinline__ float sigmoid_exact(float x) { return 1.f / (1.f + expf(-x)); }
__device__ __forceinline__ float softplus(float x) { return fmaxf(x, 0.f) + log1pf(expf(-fabsf(x))); }
// nn.BCEWithLogitsLoss(pos_weight=pw) element: (1-t) x + (1 + (pw-1) t) softplus(-x), and its derivative in x
__device__ __forceinline__ float bce_logits(float x, float t, float pw) {
  return (1.f - t) * x + (1.f + (pw - 1.f) * t) * softplus(-x);
}
__device__ __forceinline__ float bce_logits_dx(float x, float t, float pw) {
  return (1.f - t) - (1.f + (pw - 1.f) * t) * (1.f - sigmoid_exact(x));
}

__device__ __forceinline__ float block_sum(float v, float* sm) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) sm[wid] = v;
  __syncthreads();
  float r = 0.f;
  if (wid == 0) {
    r = lane < (int)(blockDim.x >> 5) ? sm[lane] : 0.f;
    r = warp_sum(r);
  }
  __syncthreads();
  return r;  // valid in warp 0
}

// ---- K1: target assignment (build_targets, loss.py:195-268) ----------------------------------------------------
__global__ void loss_assign_kernel(LossGeom g, const float* __restrict__ targets, const float* __restrict__ anchors,
                                   int* __restrict__ winner, float* __restrict__ rec, int* __restrict__ counts) {
  pdl_enter();
  const long per = 5L * g.na * g.nt;
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= per * g.nl) return;
  const int lvl = (int)(i / per);
  const int id = (int)(i % per);
  const int o = id / (g.na * g.nt), a = (id / g.nt) % g.na, t = id % g.nt;
  float* r = rec + (g.ent_off[lvl] + id) * kRec;
  int cell = -1;
  const float* tg = targets + (long)t * 6;
  const float img = tg[0];
  if (img >= 0.f && img < (float)g.bs) {
    const float nx = (float)g.nx[lvl], ny = (float)g.ny[lvl];
    const float gx = tg[2] * nx, gy = tg[3] * ny, gw = tg[4] * nx, gh = tg[5] * ny;
    const float aw = anchors[(lvl * g.na + a) * 2], ah = anchors[(lvl * g.na + a) * 2 + 1];
    const float rw = gw / aw, rh = gh / ah;
    const bool match = fmaxf(fmaxf(rw, 1.f / rw), fmaxf(rh, 1.f / rh)) < g.anchor_t;
    bool take = match;
    float ox = 0.f, oy = 0.f;
    if (o == 1) { take = take && (fmodf(gx, 1.f) < 0.5f) && (gx > 1.f); ox = 0.5f; }
    if (o == 2) { take = take && (fmodf(gy, 1.f) < 0.5f) && (gy > 1.f); oy = 0.5f; }
    if (o == 3) { const float q = nx - gx; take = take && (fmodf(q, 1.f) < 0.5f) && (q > 1.f); ox = -0.5f; }
    if (o == 4) { const float q = ny - gy; take = take && (fmodf(q, 1.f) < 0.5f) && (q > 1.f); oy = -0.5f; }
    if (take) {
      int gi = (int)(gx - ox), gj = (int)(gy - oy);  // .long(): truncation
      gi = min(max(gi, 0), g.nx[lvl] - 1);
      gj = min(max(gj, 0), g.ny[lvl] - 1);
      const int b = (int)img;
      cell = ((b * g.na + a) * g.ny[lvl] + gj) * g.nx[lvl] + gi;
      r[1] = gx - (float)gi;
      r[2] = gy - (float)gj;
      r[3] = gw;
      r[4] = gh;
      r[5] = aw;
      r[6] = ah;
      r[7] = __int_as_float((int)tg[1]);
      atomicAdd(&counts[lvl], 1);
      atomicMax(&winner[g.cell_off[lvl] + cell], id);
    }
  }
  r[0] = __int_as_float(cell);
}

struct EntryEval {
  D4 eiou;
  float sx, sy, sw, sh;  // sigmoids of the four raw box logits
  long base;             // element offset of the cell's channel 0
};

__device__ __forceinline__ EntryEval eval_entry(const LossGeom& g, int lvl, int cell, const float* r) {
  const int nx = g.nx[lvl], ny = g.ny[lvl];
  const int gi = cell % nx, gj = (cell / nx) % ny, a = (cell / (nx * ny)) % g.na, b = cell / (nx * ny * g.na);
  EntryEval e;
  e.base = b * g.sb[lvl] + a * g.sa[lvl] + gj * g.sy[lvl] + gi * g.sx[lvl];
  const float* p = g.p[lvl] + e.base;
  const long so = g.so[lvl];
  e.sx = sigmoid_exact(p[0]);
  e.sy = sigmoid_exact(p[so]);
  e.sw = sigmoid_exact(p[2 * so]);
  e.sh = sigmoid_exact(p[3 * so]);
  const float pw = (e.sw * 2.f) * (e.sw * 2.f) * r[5], ph = (e.sh * 2.f) * (e.sh * 2.f) * r[6];
  e.eiou = eiou_xywh(dvar(e.sx * 2.f - 0.5f, 0), dvar(e.sy * 2.f - 0.5f, 1), dvar(pw, 2), dvar(ph, 3), r[1], r[2], r[3],
                     r[4]);
  return e;
}

// Block partials meet in 64-bit fixed point (2^-28 units): integer addition is associative, so the loss value does not
// depend on the order in which blocks finish (fp32 atomics made the first loss of two identical runs differ by an ulp).
constexpr double kAccScale = 268435456.0;
__device__ __forceinline__ void acc_add(long long* acc, float v) {
  atomicAdd(reinterpret_cast<unsigned long long*>(acc), (unsigned long long)__double2ll_rn((double)v * kAccScale));
}
__device__ __forceinline__ float acc_get(const long long* acc) { return (float)((double)acc[0] * (1.0 / kAccScale)); }

// ---- K2: per-entry forward: box term, objectness correction on the winning entry, class term -------------------
// acc: [nl][4] = Σ(1 - eiou) | Σ softplus(x_obj) over all cells | Σ objectness correction | Σ class BCE
__global__ void __launch_bounds__(256) loss_entry_fwd_kernel(LossGeom g, const int* __restrict__ winner,
                                                             const float* __restrict__ rec, long long* __restrict__ acc) {
  pdl_enter();
  __shared__ float sm[8];
  const long per = 5L * g.na * g.nt;
  const int lvl = blockIdx.y;
  float abox = 0.f, aobj = 0.f, acls = 0.f;
  for (long id = (long)blockIdx.x * blockDim.x + threadIdx.x; id < per; id += (long)gridDim.x * blockDim.x) {
    const float* r = rec + (g.ent_off[lvl] + id) * kRec;
    const int cell = __float_as_int(r[0]);
    if (cell < 0) continue;
    const EntryEval e = eval_entry(g, lvl, cell, r);
    abox += 1.f - e.eiou.v;
    const float* p = g.p[lvl] + e.base;
    const long so = g.so[lvl];
    if (winner[g.cell_off[lvl] + cell] == (int)id) {
      const float x = p[4 * so], t = fmaxf(e.eiou.v, 0.f);
      aobj += bce_logits(x, t, g.obj_pw) - softplus(x);
    }
    if (g.nc > 1) {
      const int cls = __float_as_int(r[7]);
      for (int c = 0; c < g.nc; ++c) acls += bce_logits(p[(5 + c) * so], c == cls ? g.cp : g.cn, g.cls_pw);
    }
  }
  abox = block_sum(abox, sm);
  aobj = block_sum(aobj, sm);
  acls = block_sum(acls, sm);
  if (threadIdx.x == 0) {
    acc_add(&acc[lvl * 4 + 0], abox);
    acc_add(&acc[lvl * 4 + 2], aobj);
    acc_add(&acc[lvl * 4 + 3], acls);
  }
}

// ---- K3: dense objectness, target 0 on every cell ---------------------------------------------------------------
__global__ void __launch_bounds__(256) loss_dense_fwd_kernel(LossGeom g, long long* __restrict__ acc) {
  pdl_enter();
  __shared__ float sm[8];
  const int lvl = blockIdx.y;
  const int nx = g.nx[lvl], ny = g.ny[lvl];
  const long cells = (long)g.bs * g.na * ny * nx;
  const float* p = g.p[lvl] + 4 * g.so[lvl];
  float a = 0.f;
  for (long c = (long)blockIdx.x * blockDim.x + threadIdx.x; c < cells; c += (long)gridDim.x * blockDim.x) {
    const int gi = (int)(c % nx), gj = (int)((c / nx) % ny), an = (int)((c / ((long)nx * ny)) % g.na);
    const long b = c / ((long)nx * ny * g.na);
    a += softplus(p[b * g.sb[lvl] + an * g.sa[lvl] + gj * g.sy[lvl] + gi * g.sx[lvl]]);
  }
  a = block_sum(a, sm);
  if (threadIdx.x == 0) acc_add(&acc[lvl * 4 + 1], a);
}

// ---- K4: (lbox + lobj + lcls) * bs and the three weighted terms (loss.py:183-191) ------------------------------
__global__ void loss_finalize_kernel(LossGeom g, const long long* __restrict__ acc, const int* __restrict__ counts,
                                     float* __restrict__ out) {
  pdl_enter();
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  float lbox = 0.f, lobj = 0.f, lcls = 0.f;
  for (int l = 0; l < g.nl; ++l) {
    const float n = (float)counts[l];
    const float cells = (float)g.bs * g.na * g.ny[l] * g.nx[l];
    if (n > 0.f) {
      lbox += acc_get(&acc[l * 4 + 0]) / n;
      if (g.nc > 1) lcls += acc_get(&acc[l * 4 + 3]) / (n * g.nc);
    }
    lobj += g.balance[l] * (acc_get(&acc[l * 4 + 1]) + acc_get(&acc[l * 4 + 2])) / cells;
  }
  lbox *= g.box;
  lobj *= g.obj;
  lcls *= g.cls;
  out[0] = (lbox + lobj + lcls) * g.bs;
  out[1] = lbox;
  out[2] = lobj;
  out[3] = lcls;
}

// ---- B1: dense gradient: zero everywhere except the objectness logit ----------------------------------------------
__global__ void __launch_bounds__(256) loss_dense_bwd_kernel(LossGeom g, const float* __restrict__ gscale) {
  pdl_enter();
  const int lvl = blockIdx.y;
  const int nx = g.nx[lvl], ny = g.ny[lvl];
  const long cells = (long)g.bs * g.na * ny * nx;
  const float k = gscale[0] * g.bs * g.obj * g.balance[lvl] / (float)cells;
  const long total = cells * g.no;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int ch = (int)(i % g.no);
    const long c = i / g.no;
    const int gi = (int)(c % nx), gj = (int)((c / nx) % ny), an = (int)((c / ((long)nx * ny)) % g.na);
    const long b = c / ((long)nx * ny * g.na);
    const long off = b * g.sb[lvl] + an * g.sa[lvl] + gj * g.sy[lvl] + gi * g.sx[lvl] + ch * g.so[lvl];
    g.dp[lvl][off] = ch == 4 ? k * sigmoid_exact(g.p[lvl][off]) : 0.f;
  }
}

// ---- B2: per-entry gradient (box term through the decode, objectness correction, class term) ------------------
__global__ void __launch_bounds__(256) loss_entry_bwd_kernel(LossGeom g, const int* __restrict__ winner,
                                                             const float* __restrict__ rec,
                                                             const int* __restrict__ counts,
                                                             const float* __restrict__ gscale) {
  pdl_enter();
  const long per = 5L * g.na * g.nt;
  const int lvl = blockIdx.y;
  const float n = (float)counts[lvl];
  const float cells = (float)g.bs * g.na * g.ny[lvl] * g.nx[lvl];
  const float up = gscale[0] * g.bs;
  for (long id = (long)blockIdx.x * blockDim.x + threadIdx.x; id < per; id += (long)gridDim.x * blockDim.x) {
    const float* r = rec + (g.ent_off[lvl] + id) * kRec;
    const int cell = __float_as_int(r[0]);
    if (cell < 0) continue;
    const EntryEval e = eval_entry(g, lvl, cell, r);
    const long so = g.so[lvl];
    float* dp = g.dp[lvl] + e.base;
    const float* p = g.p[lvl] + e.base;
    // d(1 - eiou)/d(raw logits): x = 2σ - 0.5 -> 2σ(1-σ);  w = (2σ)² a -> 8 a σ² (1-σ)
    const float kb = -up * g.box / n;
    atomicAdd(dp, kb * e.eiou.d[0] * 2.f * e.sx * (1.f - e.sx));
    atomicAdd(dp + so, kb * e.eiou.d[1] * 2.f * e.sy * (1.f - e.sy));
    atomicAdd(dp + 2 * so, kb * e.eiou.d[2] * 8.f * r[5] * e.sw * e.sw * (1.f - e.sw));
    atomicAdd(dp + 3 * so, kb * e.eiou.d[3] * 8.f * r[6] * e.sh * e.sh * (1.f - e.sh));
    if (winner[g.cell_off[lvl] + cell] == (int)id) {
      const float x = p[4 * so], t = fmaxf(e.eiou.v, 0.f);
      atomicAdd(dp + 4 * so, up * g.obj * g.balance[lvl] / cells * (bce_logits_dx(x, t, g.obj_pw) - sigmoid_exact(x)));
    }
    if (g.nc > 1) {
      const int cls = __float_as_int(r[7]);
      const float kc = up * g.cls / (n * g.nc);
      for (int c = 0; c < g.nc; ++c)
        atomicAdd(dp + (5 + c) * so, kc * bce_logits_dx(p[(5 + c) * so], c == cls ? g.cp : g.cn, g.cls_pw));
    }
  }
}

static int geom_of(const ly_loss_desc* d, LossGeom* g) {
  LY_REQUIRE(d != nullptr, "loss: null descriptor");
  LY_REQUIRE(d->nl >= 1 && d->nl <= kMaxLevels && d->na >= 1 && d->na <= 8 && d->nc >= 1 && d->bs >= 1 && d->nt >= 1,
             "loss: bad sizes nl=%d na=%d nc=%d bs=%d nt=%d", d->nl, d->na, d->nc, d->bs, d->nt);
  g->nl = d->nl; g->na = d->na; g->nc = d->nc; g->no = d->nc + 5; g->bs = d->bs; g->nt = d->nt;
  long co = 0, eo = 0;
  for (int l = 0; l < d->nl; ++l) {
    LY_REQUIRE(d->p[l] != nullptr && d->ny[l] > 0 && d->nx[l] > 0, "loss: level %d has no map", l);
    g->ny[l] = d->ny[l]; g->nx[l] = d->nx[l];
    g->p[l] = d->p[l]; g->dp[l] = d->dp[l];
    g->sb[l] = d->sb[l]; g->sa[l] = d->sa[l]; g->sy[l] = d->sy[l]; g->sx[l] = d->sx[l]; g->so[l] = d->so[l];
    g->balance[l] = d->balance[l];
    g->cell_off[l] = co;
    g->ent_off[l] = eo;
    co += (long)d->bs * d->na * d->ny[l] * d->nx[l];
    eo += 5L * d->na * d->nt;
    LY_REQUIRE(co < (1L << 31), "loss: too many cells");
  }
  g->anchor_t = d->anchor_t; g->box = d->box; g->obj = d->obj; g->cls = d->cls;
  g->cls_pw = d->cls_pw; g->obj_pw = d->obj_pw; g->cp = d->cp; g->cn = d->cn;
  return LY_OK;
}

}  // namespace ly

using namespace ly;

extern "C" long ly_yolo_loss_workspace_bytes(const ly_loss_desc* d) {
  LossGeom g;
  if (geom_of(d, &g) != LY_OK) return -1;
  long cells = 0;
  for (int l = 0; l < g.nl; ++l) cells += (long)g.bs * g.na * g.ny[l] * g.nx[l];
  const long entries = 5L * g.na * g.nt * g.nl;
  // winner map | entry records | counts [4] | pad to 8 | acc [nl][4] int64
  return cells * 4 + entries * kRec * 4 + 16 + 8 + kMaxLevels * 32;
}

static void ws_split(const LossGeom& g, void* ws, int** winner, float** rec, int** counts, long long** acc, long* cells) {
  long c = 0;
  for (int l = 0; l < g.nl; ++l) c += (long)g.bs * g.na * g.ny[l] * g.nx[l];
  const long entries = 5L * g.na * g.nt * g.nl;
  *cells = c;
  *winner = (int*)ws;
  *rec = (float*)((char*)ws + c * 4);
  *counts = (int*)((char*)ws + c * 4 + entries * kRec * 4);
  *acc = (long long*)((char*)ws + ((c * 4 + entries * kRec * 4 + 16 + 7) & ~7L));
}

extern "C" int ly_yolo_loss_fwd(const ly_loss_desc* d, void* workspace, float* out, ly_stream_t st) {
  LossGeom g;
  int rc = geom_of(d, &g);
  if (rc) return rc;
  LY_REQUIRE(workspace && out && d->targets && d->anchors, "ly_yolo_loss_fwd: null pointer");
  int *winner, *counts;
  float* rec;
  long long* acc;
  long cells;
  ws_split(g, workspace, &winner, &rec, &counts, &acc, &cells);
  cudaStream_t s = (cudaStream_t)st;
  LY_CUDA(cudaMemsetAsync(winner, 0xFF, cells * 4, s));
  LY_CUDA(cudaMemsetAsync(counts, 0, 16 + 8 + kMaxLevels * 32, s));
  const long per = 5L * g.na * g.nt;
  LY_CUDA(launch(loss_assign_kernel, ceil_div(per * g.nl, 256), 256, 0, s, g, d->targets, d->anchors, winner, rec, counts));
  LY_LAUNCH_CHECK();
  dim3 ge((unsigned)min(ceil_div(per, 256), 64), (unsigned)g.nl);
  LY_CUDA(launch(loss_entry_fwd_kernel, ge, 256, 0, s, g, winner, rec, acc));
  LY_LAUNCH_CHECK();
  dim3 gd((unsigned)(device_sm_count() * 2), (unsigned)g.nl);
  LY_CUDA(launch(loss_dense_fwd_kernel, gd, 256, 0, s, g, acc));
  LY_LAUNCH_CHECK();
  LY_CUDA(launch(loss_finalize_kernel, 1, 32, 0, s, g, acc, counts, out));
  LY_LAUNCH_CHECK();
  return LY_OK;
}

extern "C" int ly_yolo_loss_bwd(const ly_loss_desc* d, const void* workspace, const float* grad_scale, ly_stream_t st) {
  LossGeom g;
  int rc = geom_of(d, &g);
  if (rc) return rc;
  LY_REQUIRE(workspace && grad_scale, "ly_yolo_loss_bwd: null pointer");
  for (int l = 0; l < g.nl; ++l) LY_REQUIRE(g.dp[l] != nullptr, "ly_yolo_loss_bwd: level %d has no gradient map", l);
  int *winner, *counts;
  float* rec;
  long long* acc;
  long cells;
  ws_split(g, const_cast<void*>(workspace), &winner, &rec, &counts, &acc, &cells);
  cudaStream_t s = (cudaStream_t)st;
  dim3 gd((unsigned)(device_sm_count() * 4), (unsigned)g.nl);
  LY_CUDA(launch(loss_dense_bwd_kernel, gd, 256, 0, s, g, grad_scale));
  LY_LAUNCH_CHECK();
  const long per = 5L * g.na * g.nt;
  dim3 ge((unsigned)min(ceil_div(per, 256), 64), (unsigned)g.nl);
  LY_CUDA(launch(loss_entry_bwd_kernel, ge, 256, 0, s, g, winner, rec, counts, grad_scale));
  LY_LAUNCH_CHECK();
  return LY_OK;
}
