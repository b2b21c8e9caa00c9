// engine.cu — host side of librpe_b200.so: the C ABI of include/rpe.h, workspace management,
// pyramid geometry / tap tables, and the launch sequence of the batched hot path
//   H2D -> pyramid -> FAST+NMS -> retainBest/Harris -> blur -> orientation+rBRIEF
//       -> Hamming k-NN -> cross-check|ratio + stable sort + cut -> 5-pt RANSAC -> recoverPose -> D2H
// on one CUDA stream, with no host synchronisation between stages.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "rpe_device.cuh"

using namespace rpe;

static thread_local std::string g_last_error;

#define RPE_CK(call)                                                                          \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess) {                                                                  \
      char buf_[512];                                                                         \
      snprintf(buf_, sizeof buf_, "%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      g_last_error = buf_;                                                                    \
      return RPE_ERR_CUDA;                                                                    \
    }                                                                                         \
  } while (0)

static int fail(int code, const char* msg) {
  g_last_error = msg;
  return code;
}
int rpe::raise_dynamic_smem_limit(const void* func, int bytes) {
  static std::mutex mu;
  static std::map<std::pair<int, const void*>, int> limit;   // (device, function) -> bytes set so far
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 1;
  std::lock_guard<std::mutex> lock(mu);
  int& cur = limit[std::make_pair(dev, func)];
  if (bytes <= cur) return 0;
  if (cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess) return 1;
  cur = bytes;
  return 0;
}

// for the entry points that live beside their kernels (kernels_eval.cu): same thread-local message as above
int rpe::report_error(int code, const char* msg) { return fail(code, msg); }

static inline int align_up(int v, int a) { return (v + a - 1) / a * a; }
static inline int pow2_ge(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

struct Geometry {
  OrbGeom g;
  std::vector<int2> taps;   // host copy
  int2* d_taps = nullptr;
  std::vector<uint32_t> ftab, btab;
  uint32_t* d_ftab = nullptr;
  uint32_t* d_btab = nullptr;
};

struct rpe_handle {
  rpe_config cfg;
  int device = 0;
  cudaStream_t own_stream = nullptr;
  cudaStream_t stream = nullptr;
  // Three-stream software pipeline, depth two batches.  The dense kernels (ORB, matcher) of every chunk
  // of every batch run in order on `stream`; image uploads run ahead on s_copy (as soon as the ORB of
  // the previous batch has released the pyramid slabs); the latency-bound geometry tail (RANSAC,
  // recoverPose, D2H; one CTA per pair) runs on the high-priority s_tail under the dense kernels of the
  // next chunk / batch.  Tail buffers exist twice (set = seq & 1).
  static constexpr int kMaxWorkers = 4;
  static constexpr int kDepth = 2;
  cudaStream_t s_copy = nullptr, s_tail = nullptr;
  // outputs of submitted, not yet waited-for batches (rpe_submit_* / rpe_wait_batch), oldest first
  struct Pending { int B = 0; int set = 0; double* R = nullptr; double* t = nullptr; double* euler = nullptr;
                   int32_t* status = nullptr; int32_t* info = nullptr; } pend[kDepth];
  int n_pending = 0;
  unsigned long long seq = 0;          // batches enqueued so far
  static constexpr int kChunkEvents = 64;
  cudaEvent_t ev_up[kChunkEvents] = {}, ev_dense[kChunkEvents] = {};
  cudaEvent_t ev_done_call[kDepth] = {};
  uint8_t* pyr2[kDepth] = {};          // the pyramid slabs exist twice too: batch n+1 is uploaded while batch n is in ORB
  // level 0 of a batch whose device-resident input already has the pyramid's layout: the caller's buffer itself
  const uint8_t* lvl0[kDepth] = {};
  size_t lvl0_stride[kDepth] = {};
  int chunk_pairs = 32, n_workers = 4;
  int NI = 0, NP = 0;
  int kp_stride = 0, cand_stride = 0, mcap = 0, sort_cap_orb = 0, sort_cap_match = 0;
  size_t pyr_stride = 0;
  OrbBuffers ob{};
  MatchBuffers mb[kDepth] = {};
  GeomBuffers gb[kDepth] = {};
  double* dK[kDepth] = {};
  int last_set = 0;          // tail-buffer set of the last completed batch / stage call (rpe_get_stage)
  std::vector<void*> allocs;
  std::map<std::pair<int, int>, Geometry> geoms;
  const Geometry* last_geom = nullptr;
  int last_pairs = 0;
  // input format of the batch / sequence calls (rpe_set_input_format) + lazily allocated BGR staging
  int in_channels = 1, gray_mode = RPE_GRAY_CVTCOLOR;
  double* eval_buf = nullptr;   // rpe_evaluate_sequence scratch (grown on demand)
  size_t eval_cap = 0;
  uint8_t* bgr_stage = nullptr;
  size_t bgr_stage_stride = 0, bgr_stage_imgs = 0;
  int last_pstep = 2;       // image slots of pair p in the last batch: pstep*p, pstep*p + 1
  // pinned host staging for results: R(9) t(3) euler(3) doubles, status + 4 info ints, K
  double* h_out_d[kDepth] = {};
  int32_t* h_out_i[kDepth] = {};   // status | num_matches | pose count | rinfo (4) per pair
  double* h_K[kDepth] = {};
  // profiling
  bool profiling = false;
  cudaEvent_t ev[RPE_T_COUNT + 1] = {};
  float ms[RPE_T_COUNT] = {};
  int launches[RPE_T_COUNT] = {};
};

// ---------------------------------------------------------------------------------------------
// Pyramid geometry exactly as ORB_Impl: f32 scale factor, f32 level scales,
// cvRound level sizes, f32 per-level quotas; plus INTER_LINEAR_EXACT tap tables in f64.
static void compute_levels(const rpe_config& c, int W, int H, int* w, int* h, float* scale, int* quota) {
  const double sf = (double)c.scale_factor;
  for (int l = 0; l < c.nlevels; ++l) {
    scale[l] = (float)std::pow(sf, (double)l);
    w[l] = (int)lrintf((float)W / scale[l]);
    h[l] = (int)lrintf((float)H / scale[l]);
  }
  const float factor = (float)(1.0 / sf);
  float nd = c.nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)c.nlevels));
  int sum = 0;
  for (int l = 0; l < c.nlevels - 1; ++l) {
    quota[l] = (int)lrintf(nd);
    sum += quota[l];
    nd *= factor;
  }
  quota[c.nlevels - 1] = std::max(c.nfeatures - sum, 0);
}

static void linear_exact_taps(int src, int dst, std::vector<int2>& out) {
  const double scale = 1.0 / ((double)dst / (double)src);
  for (int d = 0; d < dst; ++d) {
    const double f = scale * (d + 0.5) - 0.5;
    int i = (int)std::floor(f);
    int c1 = (int)lrint((f - i) * 256.0);
    if (i < 0 || src <= 1) { i = 0; c1 = 0; }
    else if (i >= src - 1) { i = src - 1; c1 = 0; }
    out.push_back(make_int2(i, c1));
  }
}

static int kp_cap_of(int quota) { return quota + std::max(64, quota / 4); }
static int cand_cap_of(int w, int h) { return std::max(1024, (w * h) / 16); }

static void build_geometry(const rpe_config& c, int W, int H, Geometry& G) {
  OrbGeom& g = G.g;
  memset(&g, 0, sizeof g);
  int w[RPE_MAX_LEVELS], h[RPE_MAX_LEVELS], quota[RPE_MAX_LEVELS];
  float scale[RPE_MAX_LEVELS];
  compute_levels(c, W, H, w, h, scale, quota);
  g.nlevels = c.nlevels;
  g.fast_threshold = c.fast_threshold;
  size_t off = 0;
  int cand_off = 0, kp_off = 0, tile0 = 0, ftile0 = 0, max_kp_cap = 1;
  G.taps.clear();
  G.ftab.clear();
  G.btab.clear();
  for (int l = 0; l < c.nlevels; ++l) {
    LevelGeom& L = g.lv[l];
    L.w = w[l]; L.h = h[l]; L.pitch = align_up(std::max(w[l], 1), 16);
    L.quota = quota[l];
    L.scale = scale[l];
    L.off = (uint32_t)off;
    off += (size_t)align_up(L.pitch * std::max(L.h, 1), 256);
    L.cand_off = cand_off; L.cand_cap = cand_cap_of(L.w, L.h); cand_off += L.cand_cap;
    L.kp_off = kp_off; L.kp_cap = kp_cap_of(L.quota); kp_off += L.kp_cap;
    max_kp_cap = std::max(max_kp_cap, L.kp_cap);
    L.tiles_x = (L.w + kTileW - 1) / kTileW; L.tiles_y = (L.h + kTileH - 1) / kTileH;
    L.tile0 = tile0; tile0 += L.tiles_x * L.tiles_y;
    // FAST tiles cover x, y in [kFastOrigin, size - 31): keypoints live in [31, size - 31) and their
    // NMS neighbours in [30, size - 30); levels narrower than 63 px have no tiles at all
    const int fw = L.w - 31 - kFastOrigin, fh = L.h - 31 - kFastOrigin;
    L.ftiles_x = fw > 0 && fh > 0 ? (fw + kFastTW - 1) / kFastTW : 0;
    L.ftiles_y = fw > 0 && fh > 0 ? (fh + kFastTH - 1) / kFastTH : 0;
    L.ftile0 = ftile0; ftile0 += L.ftiles_x * L.ftiles_y;
    for (int ty = 0; ty < L.ftiles_y; ++ty)
      for (int tx = 0; tx < L.ftiles_x; ++tx) G.ftab.push_back(((uint32_t)l << 24) | ((uint32_t)ty << 12) | (uint32_t)tx);
    for (int sy = 0; sy < (L.h + 2 * kBlurRows - 1) / (2 * kBlurRows); ++sy)
      for (int sx = 0; sx < (L.w + 127) / 128; ++sx) G.btab.push_back(((uint32_t)l << 24) | ((uint32_t)sy << 12) | (uint32_t)sx);
    if (l > 0) {
      L.xtab = (int)G.taps.size(); linear_exact_taps(w[l - 1], w[l], G.taps);
      L.ytab = (int)G.taps.size(); linear_exact_taps(h[l - 1], h[l], G.taps);
    }
  }
  // k_resize_strip needs x0[k] - x0[first of its aligned group of 4] <= 6 for every output column
  g.resize_strip_ok = 1;
  for (int l = 1; l < c.nlevels; ++l) {
    const int2* xt = G.taps.data() + g.lv[l].xtab;
    for (int x = 0; x < w[l]; ++x)
      if (xt[x].x - xt[x & ~3].x > 6) g.resize_strip_ok = 0;
  }
  // source rows an 8-row output strip needs: rows y0[first] .. y0[last] + 1
  g.resize_span = 0;
  for (int l = 1; l < c.nlevels; ++l) {
    const int2* yt = G.taps.data() + g.lv[l].ytab;
    for (int y = 0; y < h[l]; y += 8)
      g.resize_span = std::max(g.resize_span, yt[std::min(y + 7, h[l] - 1)].x + 2 - yt[y].x);
  }
  // source bytes a group of 128 output columns needs, counted from the 16-byte-aligned column at or left of its first
  // tap: the last lane (4 columns from x & ~3 <= w - 1) reads three words from its own aligned tap
  g.resize_rows_strict = 1;
  for (int l = 1; l < c.nlevels; ++l) {
    const int2* yt = G.taps.data() + g.lv[l].ytab;
    for (int y = 1; y < h[l]; ++y)
      if (yt[y].x <= yt[y - 1].x) g.resize_rows_strict = 0;
  }
  g.resize_xspan = 0;
  for (int l = 1; l < c.nlevels; ++l) {
    const int2* xt = G.taps.data() + g.lv[l].xtab;
    for (int x = 0; x < g.lv[l].pitch; x += 128) {
      const int astart = xt[std::min(x, w[l] - 1)].x & ~15;
      for (int q = x; q <= std::min(x + 124, g.lv[l].pitch - 4); q += 4)
        g.resize_xspan = std::max(g.resize_xspan, (xt[std::min(q, w[l] - 1)].x & ~3) + 12 - astart);
    }
  }
  g.total_tiles = tile0;
  g.total_ftiles = ftile0;
  g.total_bitems = (int)G.btab.size();
  g.cand_stride = cand_off;
  g.kp_stride = kp_off;
  g.sort_cap = align_up(max_kp_cap, 8);   // (no power of two needed: k_select orders by rows, not by a bitonic network)
  g.sel_row_cap = H;
  g.pyr_stride = off;
}

template <typename T>
static int dalloc(rpe_handle* h, T** p, size_t n) {
  void* q = nullptr;
  RPE_CK(cudaMalloc(&q, std::max<size_t>(n, 1) * sizeof(T)));
  RPE_CK(cudaMemset(q, 0, std::max<size_t>(n, 1) * sizeof(T)));
  h->allocs.push_back(q);
  *p = (T*)q;
  return RPE_OK;
}
#define RPE_TRY(x)            \
  do {                        \
    int r_ = (x);             \
    if (r_ != RPE_OK) return r_; \
  } while (0)

static int get_geometry(rpe_handle* h, int W, int H, const Geometry** out) {
  auto key = std::make_pair(W, H);
  auto it = h->geoms.find(key);
  if (it == h->geoms.end()) {
    Geometry G;
    build_geometry(h->cfg, W, H, G);
    // strides are those of the handle (max image), so every slot has the same footprint
    if (G.g.pyr_stride > h->pyr_stride || G.g.cand_stride > h->cand_stride) return fail(RPE_ERR_CAPACITY, "image larger than handle capacity");
    G.g.pyr_stride = h->pyr_stride;
    G.g.cand_stride = h->cand_stride;
    G.g.kp_stride = h->kp_stride;
    G.g.sort_cap = h->sort_cap_orb;
    G.g.sel_row_cap = h->cfg.max_height;
    RPE_CK(cudaMalloc((void**)&G.d_taps, std::max<size_t>(G.taps.size(), 1) * sizeof(int2)));
    RPE_CK(cudaMemcpy(G.d_taps, G.taps.data(), G.taps.size() * sizeof(int2), cudaMemcpyHostToDevice));
    RPE_CK(cudaMalloc((void**)&G.d_ftab, std::max<size_t>(G.ftab.size(), 1) * sizeof(uint32_t)));
    RPE_CK(cudaMemcpy(G.d_ftab, G.ftab.data(), G.ftab.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    RPE_CK(cudaMalloc((void**)&G.d_btab, std::max<size_t>(G.btab.size(), 1) * sizeof(uint32_t)));
    RPE_CK(cudaMemcpy(G.d_btab, G.btab.data(), G.btab.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    it = h->geoms.emplace(key, std::move(G)).first;
  }
  *out = &it->second;
  return RPE_OK;
}

extern "C" {

int rpe_abi_version(void) { return RPE_ABI_VERSION; }
const char* rpe_last_error(void) { return g_last_error.c_str(); }

void rpe_default_config(rpe_config* c) {
  memset(c, 0, sizeof *c);
  c->max_width = 640; c->max_height = 480; c->max_pairs = 1;
  c->nfeatures = 4000; c->nlevels = 12; c->scale_factor = 1.1f; c->fast_threshold = 15;
  c->match_mode = RPE_MATCH_CROSSCHECK; c->ratio = 0.75; c->max_matches = 500;
  c->ransac_prob = 0.999; c->ransac_threshold = 1.0; c->ransac_max_iters = 1000;
  c->euler_convention = RPE_EULER_YUP;
}

int rpe_create(const rpe_config* cfg, int device, rpe_handle** out) {
  if (!cfg || !out) return fail(RPE_ERR_INVALID, "null argument");
  *out = nullptr;
  if (cfg->nlevels < 1 || cfg->nlevels > RPE_MAX_LEVELS || cfg->nfeatures < 1 || cfg->max_pairs < 1 ||
      cfg->max_width < 1 || cfg->max_height < 1 || cfg->max_width > RPE_MAX_DIM || cfg->max_height > RPE_MAX_DIM ||
      !(cfg->scale_factor > 1.0f) || cfg->fast_threshold < 1 || cfg->fast_threshold > 254 ||
      (cfg->match_mode != RPE_MATCH_CROSSCHECK && cfg->match_mode != RPE_MATCH_RATIO) || cfg->ransac_max_iters < 1)
    return fail(RPE_ERR_INVALID, "invalid rpe_config");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0 || device < 0 || device >= ndev)
    return fail(RPE_ERR_NO_DEVICE, "no usable CUDA device (this library has no CPU fallback)");
  cudaDeviceProp prop;
  RPE_CK(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) return fail(RPE_ERR_NO_DEVICE, "device is not sm_100 (B200); kernels are built for sm_100a only");
  RPE_CK(cudaSetDevice(device));
  rpe_handle* h = new rpe_handle();
  h->cfg = *cfg;
  h->device = device;
  h->NP = cfg->max_pairs;
  h->NI = 2 * cfg->max_pairs;
  Geometry G;
  build_geometry(*cfg, cfg->max_width, cfg->max_height, G);
  // a narrower-but-taller image can need more pyramid bytes through pitch rounding: pad a little
  h->pyr_stride = G.g.pyr_stride + 4096;
  h->cand_stride = G.g.cand_stride + 1024 * cfg->nlevels;
  h->kp_stride = G.g.kp_stride;
  h->sort_cap_orb = G.g.sort_cap;
  if (h->kp_stride >= (1 << 20)) { delete h; return fail(RPE_ERR_INVALID, "nfeatures too large (index packs in 20 bits)"); }
  h->mcap = cfg->max_matches > 0 ? std::min(cfg->max_matches, h->kp_stride) : h->kp_stride;
  h->sort_cap_match = pow2_ge(h->mcap);
  if ((size_t)h->sort_cap_match * 4 > 200 * 1024 || h->sort_cap_orb > 65536 ||
      select_smem_bytes(h->sort_cap_orb, cfg->max_height) > 200 * 1024) {
    delete h;
    return fail(RPE_ERR_INVALID, "max_matches / nfeatures exceed the shared-memory sort capacity");
  }
  *out = h;   // so that rpe_destroy can clean up on failure below
  int rc = RPE_OK;
  auto A = [&](int r) { if (rc == RPE_OK) rc = r; };
  const size_t NI = h->NI, NP = h->NP, L = cfg->nlevels, KS = h->kp_stride, CS = h->cand_stride, MC = h->mcap;
  for (int k = 0; k < rpe_handle::kDepth; ++k) A(dalloc(h, &h->pyr2[k], NI * h->pyr_stride));
  h->ob.pyr = h->pyr2[0];
  A(dalloc(h, &h->ob.blur, NI * h->pyr_stride));
  A(dalloc(h, &h->ob.cand_key, NI * CS));
  // cand_cnt, hist and img_flags are cleared with one memset per batch: keep them contiguous
  uint32_t* ctr = nullptr;
  A(dalloc(h, &ctr, NI * L + NI * L * 256 + NI));
  h->ob.cand_cnt = ctr; h->ob.hist = ctr + NI * L; h->ob.img_flags = ctr + NI * L + NI * L * 256;
  A(dalloc(h, &h->ob.sel_key, NI * CS));
  A(dalloc(h, &h->ob.sel_resp, NI * CS));
  A(dalloc(h, &h->ob.lvl_key, NI * KS));
  A(dalloc(h, &h->ob.lvl_resp, NI * KS));
  A(dalloc(h, &h->ob.lvl_cnt, NI * L));
  A(dalloc(h, &h->ob.kp_x, NI * KS)); A(dalloc(h, &h->ob.kp_y, NI * KS)); A(dalloc(h, &h->ob.kp_size, NI * KS));
  A(dalloc(h, &h->ob.kp_angle, NI * KS)); A(dalloc(h, &h->ob.kp_resp, NI * KS)); A(dalloc(h, &h->ob.kp_oct, NI * KS));
  A(dalloc(h, &h->ob.desc, NI * KS * 32));
  A(dalloc(h, &h->ob.kp_n, NI));
  A(dalloc(h, &h->mb[0].fwd1, NP * KS)); A(dalloc(h, &h->mb[0].fwd2, NP * KS)); A(dalloc(h, &h->mb[0].bwd1, NP * KS));
  for (int k = 0; k < rpe_handle::kDepth; ++k) {
    MatchBuffers& mb = h->mb[k];
    GeomBuffers& gb = h->gb[k];
    mb.fwd1 = h->mb[0].fwd1; mb.fwd2 = h->mb[0].fwd2; mb.bwd1 = h->mb[0].bwd1;   // dense-stream only: one copy
    A(dalloc(h, &mb.m_q, NP * MC)); A(dalloc(h, &mb.m_t, NP * MC)); A(dalloc(h, &mb.m_d, NP * MC));
    A(dalloc(h, &mb.m_n, NP)); A(dalloc(h, &mb.pstat, NP));
    A(dalloc(h, &mb.pts1, NP * MC * 2)); A(dalloc(h, &mb.pts2, NP * MC * 2));
    A(dalloc(h, &gb.npts, NP * MC * 4));
    A(dalloc(h, &gb.E, NP * 9));
    A(dalloc(h, &gb.rmask, NP * MC)); A(dalloc(h, &gb.pmask, NP * MC));
    A(dalloc(h, &gb.rinfo, NP * 4));
    A(dalloc(h, &gb.rstate, NP)); A(dalloc(h, &gb.rsubsets, NP * kRansacSuperRoundCap * 5));
    A(dalloc(h, &gb.rcands, NP * kRansacCandPerPair)); A(dalloc(h, &gb.rncand, NP * 8));
    A(dalloc(h, &gb.rmodels, NP * kRansacSuperRoundCap * 90)); A(dalloc(h, &gb.rnm, NP * kRansacSuperRoundCap));
    A(dalloc(h, &gb.R, NP * 9)); A(dalloc(h, &gb.t, NP * 3)); A(dalloc(h, &gb.euler, NP * 3));
    A(dalloc(h, &gb.pose_cnt, NP)); A(dalloc(h, &gb.status, NP));
    A(dalloc(h, &h->dK[k], 9));
  }
  if (rc == RPE_OK && cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking) != cudaSuccess) rc = fail(RPE_ERR_CUDA, "cudaStreamCreate failed");
  h->stream = h->own_stream;
  if (rc == RPE_OK) {
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    if (cudaStreamCreateWithFlags(&h->s_copy, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithPriority(&h->s_tail, cudaStreamNonBlocking, prio_hi) != cudaSuccess)
      rc = fail(RPE_ERR_CUDA, "pipeline stream creation failed");
  }
  for (int i = 0; i < rpe_handle::kChunkEvents && rc == RPE_OK; ++i)
    if (cudaEventCreateWithFlags(&h->ev_up[i], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&h->ev_dense[i], cudaEventDisableTiming) != cudaSuccess) rc = fail(RPE_ERR_CUDA, "cudaEventCreate failed");
  for (int k = 0; k < rpe_handle::kDepth && rc == RPE_OK; ++k) {
    if (cudaEventCreateWithFlags(&h->ev_done_call[k], cudaEventDisableTiming) != cudaSuccess) rc = fail(RPE_ERR_CUDA, "cudaEventCreate failed");
    if (rc == RPE_OK && (cudaMallocHost((void**)&h->h_out_d[k], NP * 15 * sizeof(double)) != cudaSuccess ||
                         cudaMallocHost((void**)&h->h_out_i[k], NP * 7 * sizeof(int32_t)) != cudaSuccess ||
                         cudaMallocHost((void**)&h->h_K[k], 9 * sizeof(double)) != cudaSuccess))
      rc = fail(RPE_ERR_CUDA, "cudaMallocHost failed");
  }
  if (rc == RPE_OK) {
    for (int i = 0; i <= RPE_T_COUNT; ++i)
      if (cudaEventCreate(&h->ev[i]) != cudaSuccess) rc = fail(RPE_ERR_CUDA, "cudaEventCreate failed");
  }
  if (rc == RPE_OK && (setup_orb_attributes((int)select_smem_bytes(h->sort_cap_orb, h->cfg.max_height)) || setup_match_attributes(h->sort_cap_match) || setup_geom_attributes()))
    rc = fail(RPE_ERR_CUDA, "cudaFuncSetAttribute failed");
  if (rc != RPE_OK) {
    std::string keep = g_last_error;
    rpe_destroy(h);
    g_last_error = keep;
    *out = nullptr;
  }
  return rc;
}

void rpe_destroy(rpe_handle* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  for (void* p : h->allocs) cudaFree(p);
  for (auto& kv : h->geoms) { cudaFree(kv.second.d_taps); cudaFree(kv.second.d_ftab); cudaFree(kv.second.d_btab); }
  if (h->bgr_stage) cudaFree(h->bgr_stage);
  if (h->eval_buf) cudaFree(h->eval_buf);
  if (h->gb[0].tr_subset) { cudaFree(h->gb[0].tr_subset); cudaFree(h->gb[0].tr_nmodels); cudaFree(h->gb[0].tr_counts); cudaFree(h->gb[0].tr_models); }
  for (int k = 0; k < rpe_handle::kDepth; ++k) {
    if (h->h_out_d[k]) cudaFreeHost(h->h_out_d[k]);
    if (h->h_out_i[k]) cudaFreeHost(h->h_out_i[k]);
    if (h->h_K[k]) cudaFreeHost(h->h_K[k]);
    if (h->ev_done_call[k]) cudaEventDestroy(h->ev_done_call[k]);
  }
  for (int i = 0; i <= RPE_T_COUNT; ++i) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
  if (h->own_stream) cudaStreamDestroy(h->own_stream);
  if (h->s_copy) cudaStreamDestroy(h->s_copy);
  if (h->s_tail) cudaStreamDestroy(h->s_tail);
  for (int i = 0; i < rpe_handle::kChunkEvents; ++i) {
    if (h->ev_up[i]) cudaEventDestroy(h->ev_up[i]);
    if (h->ev_dense[i]) cudaEventDestroy(h->ev_dense[i]);
  }
  delete h;
}

int rpe_set_stream(rpe_handle* h, void* s) {
  if (!h) return fail(RPE_ERR_INVALID, "null handle");
  if (h->n_pending) return fail(RPE_ERR_INVALID, "a submitted batch is pending: call rpe_wait_batch first");
  h->stream = s ? (cudaStream_t)s : h->own_stream;
  return RPE_OK;
}

int rpe_set_pipeline(rpe_handle* h, int32_t chunk_pairs, int32_t n_streams) {
  if (!h || chunk_pairs < 1 || n_streams < 1 || n_streams > rpe_handle::kMaxWorkers) return fail(RPE_ERR_INVALID, "bad pipeline arguments");
  h->chunk_pairs = chunk_pairs;
  h->n_workers = n_streams;
  return RPE_OK;
}

int rpe_set_input_format(rpe_handle* h, int32_t channels, int32_t gray_mode) {
  if (!h || (channels != 1 && channels != 3) || (gray_mode != RPE_GRAY_CVTCOLOR && gray_mode != RPE_GRAY_IMREAD_PNG))
    return fail(RPE_ERR_INVALID, "channels must be 1 or 3, gray_mode RPE_GRAY_*");
  if (h->n_pending) return fail(RPE_ERR_INVALID, "a submitted batch is pending: call rpe_wait_batch first");
  h->in_channels = channels;
  h->gray_mode = gray_mode;
  return RPE_OK;
}

int rpe_get_capacity(const rpe_handle* h, int32_t* kp_cap, int32_t* match_cap) {
  if (!h) return fail(RPE_ERR_INVALID, "null handle");
  if (kp_cap) *kp_cap = h->kp_stride;
  if (match_cap) *match_cap = h->mcap;
  return RPE_OK;
}

int rpe_set_profiling(rpe_handle* h, int32_t on) {
  if (!h) return fail(RPE_ERR_INVALID, "null handle");
  h->profiling = on != 0;
  return RPE_OK;
}

int rpe_get_stage_times(rpe_handle* h, float* ms, int32_t* launches) {
  if (!h) return fail(RPE_ERR_INVALID, "null handle");
  for (int i = 0; i < RPE_T_COUNT; ++i) {
    if (ms) ms[i] = h->ms[i];
    if (launches) launches[i] = h->launches[i];
  }
  return RPE_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Slot views: the same buffer structs with every pointer advanced to image slot img0 = 2 * pair0,
// so the kernels (which index from 0) work on a sub-batch.
static OrbBuffers orb_view(const rpe_handle* h, int set, int img0) {
  OrbBuffers b = h->ob;
  b.pyr = h->pyr2[set];
  b.lvl0 = h->lvl0[set] ? h->lvl0[set] + (size_t)img0 * h->lvl0_stride[set] : nullptr;
  b.lvl0_stride = h->lvl0_stride[set];
  const size_t i = (size_t)img0, L = h->cfg.nlevels, CS = h->cand_stride, KS = h->kp_stride;
  b.pyr += i * h->pyr_stride; b.blur += i * h->pyr_stride;
  b.cand_key += i * CS; b.sel_key += i * CS; b.sel_resp += i * CS;
  b.cand_cnt += i * L; b.hist += i * L * 256; b.lvl_cnt += i * L; b.img_flags += i;
  b.lvl_key += i * KS; b.lvl_resp += i * KS;
  b.kp_x += i * KS; b.kp_y += i * KS; b.kp_size += i * KS; b.kp_angle += i * KS; b.kp_resp += i * KS;
  b.kp_oct += i * KS; b.desc += i * KS * 32; b.kp_n += i;
  return b;
}
static MatchBuffers match_view(const rpe_handle* h, int set, int pair0) {
  MatchBuffers m = h->mb[set];
  const size_t p = (size_t)pair0, KS = h->kp_stride, MC = h->mcap;
  m.fwd1 += p * KS; m.fwd2 += p * KS; m.bwd1 += p * KS;
  m.m_q += p * MC; m.m_t += p * MC; m.m_d += p * MC; m.m_n += p; m.pstat += p;
  m.pts1 += p * MC * 2; m.pts2 += p * MC * 2;
  return m;
}
static GeomBuffers geom_view(const rpe_handle* h, int set, int pair0) {
  GeomBuffers g = h->gb[set];
  const size_t p = (size_t)pair0, MC = h->mcap;
  g.npts += p * MC * 4; g.E += p * 9; g.rmask += p * MC; g.pmask += p * MC; g.rinfo += p * 4;
  g.rstate += p; g.rsubsets += p * kRansacSuperRoundCap * 5; g.rcands += p * kRansacCandPerPair; g.rncand += p * 8;
  g.rmodels += p * kRansacSuperRoundCap * 90; g.rnm += p * kRansacSuperRoundCap;
  g.R += p * 9; g.t += p * 3; g.euler += p * 3; g.pose_cnt += p; g.status += p;
  g.tr_subset = nullptr;
  return g;
}

static inline void mark(rpe_handle* h, int i, bool prof) {
  if (prof) cudaEventRecord(h->ev[i], h->stream);   // profiling runs every stage on the handle's stream
}

static int finish_profile(rpe_handle* h) {
  if (!h->profiling) return RPE_OK;
  for (int i = 0; i < RPE_T_COUNT; ++i) RPE_CK(cudaEventElapsedTime(&h->ms[i], h->ev[i], h->ev[i + 1]));
  return RPE_OK;
}

static int clear_counters(rpe_handle* h, cudaStream_t s) {
  const size_t nctr = (size_t)h->NI * h->cfg.nlevels * 257 + h->NI;
  RPE_CK(cudaMemsetAsync(h->ob.cand_cnt, 0, nctr * sizeof(uint32_t), s));
  return RPE_OK;
}

// ORB over image slots [img0, img0 + n_img) on stream s (counters already cleared)
static int run_orb(rpe_handle* h, const Geometry& G, int set, int img0, int n_img, cudaStream_t s, bool prof) {
  OrbBuffers b = orb_view(h, set, img0);
  b.taps = G.d_taps;
  b.ftab = G.d_ftab;
  b.btab = G.d_btab;
  int* ln = h->launches;
  mark(h, RPE_T_PYRAMID, prof);
  launch_pyramid(G.g, b, n_img, s, &ln[RPE_T_PYRAMID]);
  mark(h, RPE_T_FAST, prof);
  launch_fast(G.g, b, n_img, s, &ln[RPE_T_FAST]);
  mark(h, RPE_T_SELECT, prof);
  launch_select(G.g, b, n_img, s, &ln[RPE_T_SELECT]);
  mark(h, RPE_T_BLUR, prof);
  launch_blur(G.g, b, n_img, s, &ln[RPE_T_BLUR]);
  mark(h, RPE_T_DESCRIBE, prof);
  launch_describe(G.g, b, n_img, s, &ln[RPE_T_DESCRIBE]);
  RPE_CK(cudaGetLastError());
  return RPE_OK;
}

static RansacParams ransac_params(const rpe_handle* h) {
  RansacParams rp;
  rp.prob = h->cfg.ransac_prob; rp.threshold = h->cfg.ransac_threshold;
  rp.max_iters = h->cfg.ransac_max_iters; rp.euler_convention = h->cfg.euler_convention;
  return rp;
}

// match -> sort/cut over pair slots [pair0, pair0 + nb) on stream s (dense, XU-pipe bound)
static int run_match(rpe_handle* h, int set, int pair0, int nb, int pstep, cudaStream_t s, bool prof) {
  int* ln = h->launches;
  const OrbBuffers ob = orb_view(h, set, pstep * pair0);
  const MatchBuffers mb = match_view(h, set, pair0);
  mark(h, RPE_T_MATCH, prof);
  launch_match(ob, mb, h->kp_stride, h->cfg.nfeatures, nb, pstep, h->cfg.match_mode, s, &ln[RPE_T_MATCH]);
  mark(h, RPE_T_MATCH_SORT, prof);
  launch_build_matches(ob, mb, h->kp_stride, h->mcap, h->sort_cap_match, nb, pstep, h->cfg.match_mode, h->cfg.ratio,
                       h->cfg.max_matches, s, &ln[RPE_T_MATCH_SORT]);
  RPE_CK(cudaGetLastError());
  return RPE_OK;
}

// RANSAC -> pose over pair slots [pair0, pair0 + nb) on stream s (one CTA per pair: latency bound)
// (reads only the tail-buffer set: never the ORB buffers, which the next batch may be overwriting)
static int run_geometry(rpe_handle* h, int set, int pair0, int nb, cudaStream_t s, bool prof) {
  int* ln = h->launches;
  const MatchBuffers mb = match_view(h, set, pair0);
  const GeomBuffers gb = geom_view(h, set, pair0);
  mark(h, RPE_T_RANSAC, prof);
  launch_normalize(mb, gb, h->mcap, nb, h->dK[set], s, &ln[RPE_T_RANSAC]);
  launch_ransac(mb, gb, h->mcap, nb, h->dK[set], ransac_params(h), s, &ln[RPE_T_RANSAC]);
  mark(h, RPE_T_POSE, prof);
  launch_pose(mb, gb, h->mcap, nb, ransac_params(h), mb.pstat, s, &ln[RPE_T_POSE]);
  RPE_CK(cudaGetLastError());
  return RPE_OK;
}

static int upload_K(rpe_handle* h, int set, const double K[9]) {
  memcpy(h->h_K[set], K, 9 * sizeof(double));
  RPE_CK(cudaMemcpyAsync(h->dK[set], h->h_K[set], 9 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  return RPE_OK;
}

// async D2H of the results of pair slots [pair0, pair0 + nb) of a set into its pinned staging buffers
static int download_async(rpe_handle* h, int set, int pair0, int nb, cudaStream_t s) {
  double* hd = h->h_out_d[set];
  int32_t* hi = h->h_out_i[set];
  const GeomBuffers& gb = h->gb[set];
  const size_t NP = h->NP, p = pair0, n = nb;
  RPE_CK(cudaMemcpyAsync(hd + p * 9, gb.R + p * 9, n * 9 * sizeof(double), cudaMemcpyDeviceToHost, s));
  RPE_CK(cudaMemcpyAsync(hd + NP * 9 + p * 3, gb.t + p * 3, n * 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
  RPE_CK(cudaMemcpyAsync(hd + NP * 12 + p * 3, gb.euler + p * 3, n * 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
  RPE_CK(cudaMemcpyAsync(hi + p, gb.status + p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  RPE_CK(cudaMemcpyAsync(hi + NP + p, h->mb[set].m_n + p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  RPE_CK(cudaMemcpyAsync(hi + 2 * NP + p, gb.pose_cnt + p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  RPE_CK(cudaMemcpyAsync(hi + 3 * NP + p * 4, gb.rinfo + p * 4, n * 4 * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  return RPE_OK;
}

// wait for a batch's tail (its last D2H) and hand the staged results to the caller
static int finish_results(rpe_handle* h, const rpe_handle::Pending& p) {
  const double* hd = h->h_out_d[p.set];
  const int32_t* hi = h->h_out_i[p.set];
  const size_t NP = h->NP, B = p.B;
  RPE_CK(cudaEventSynchronize(h->ev_done_call[p.set]));
  if (p.R) memcpy(p.R, hd, B * 9 * sizeof(double));
  if (p.t) memcpy(p.t, hd + NP * 9, B * 3 * sizeof(double));
  if (p.euler) memcpy(p.euler, hd + NP * 12, B * 3 * sizeof(double));
  if (p.status) memcpy(p.status, hi, B * sizeof(int32_t));
  if (p.info) {
    for (size_t i = 0; i < B; ++i) {
      p.info[4 * i + 0] = hi[NP + i];                 // num_matches
      p.info[4 * i + 1] = hi[2 * NP + i];             // recoverPose count
      p.info[4 * i + 2] = hi[3 * NP + 4 * i + 1];     // RANSAC inliers
      p.info[4 * i + 3] = hi[3 * NP + 4 * i + 2];     // RANSAC iterations
    }
  }
  h->last_set = p.set;
  return finish_profile(h);
}

static int check_batch(rpe_handle* h, int B, int H, int W, size_t pitch) {
  if (!h) return fail(RPE_ERR_INVALID, "null handle");
  if (B < 1 || H < 1 || W < 1) return fail(RPE_ERR_INVALID, "bad batch arguments");
  if (B > h->NP) return fail(RPE_ERR_CAPACITY, "batch larger than max_pairs");
  if (W > RPE_MAX_DIM || H > RPE_MAX_DIM) return fail(RPE_ERR_CAPACITY, "image dimension above 4095");
  RPE_CK(cudaSetDevice(h->device));
  return RPE_OK;
}

static void reset_profile(rpe_handle* h) {
  for (int i = 0; i < RPE_T_COUNT; ++i) { h->launches[i] = 0; h->ms[i] = 0.f; }
}

// Source of one batch: host pointer arrays or one device buffer.  Pair batch: image slot 2i = img1[i],
// 2i + 1 = img2[i], B pairs.  Sequence (frame-level feature cache): image slot i = img1[i]
// (F frames), pair p = frames (p, p + 1), so every interior frame goes through ORB once, not twice.
struct BatchSource {
  const uint8_t* const* img1 = nullptr;
  const uint8_t* const* img2 = nullptr;
  bool sequence = false;
  int channels = 1;               // 1: gray; 3: interleaved BGR, converted on the device (k_bgr2gray)
  bool host_contiguous = false;   // all host images equally spaced in slot order, device pitch
  ptrdiff_t host_stride = 0;
  const uint8_t* d_images = nullptr;
  size_t img_stride = 0;
  size_t pitch = 0;
  const uint8_t* host_image(int slot) const { return sequence ? img1[slot] : ((slot & 1) ? img2[slot >> 1] : img1[slot >> 1]); }
};

// staging buffer for interleaved BGR input: n_img slots of H rows x align16(3 W) bytes
static int ensure_bgr_stage(rpe_handle* h, int n_img, int H, int W) {
  const size_t stride = (size_t)align_up(align_up(3 * W, 16) * H, 256);
  if (h->bgr_stage && h->bgr_stage_stride >= stride && h->bgr_stage_imgs >= (size_t)n_img) return RPE_OK;
  RPE_CK(cudaDeviceSynchronize());
  if (h->bgr_stage) cudaFree(h->bgr_stage);
  h->bgr_stage = nullptr;
  const size_t imgs = std::max<size_t>(n_img, h->bgr_stage_imgs), st = std::max(stride, h->bgr_stage_stride);
  RPE_CK(cudaMalloc((void**)&h->bgr_stage, imgs * st + 64));
  h->bgr_stage_stride = st;
  h->bgr_stage_imgs = imgs;
  return RPE_OK;
}

// copy the images of slots [img0, img0 + ni) into level 0 of their pyramid slabs; BGR input goes
// through the staging buffer (or straight from an aligned device buffer) and k_bgr2gray
static int upload_images(rpe_handle* h, const Geometry& G, const BatchSource& src, int set, int img0, int ni, int H,
                         int W, cudaStream_t s) {
  const bool bgr = src.channels == 3;
  const int p0 = G.g.lv[0].pitch;                                    // gray pitch of level 0
  const int dp = bgr ? align_up(3 * W, 16) : p0;                     // device pitch of the copied rows
  const size_t row = (size_t)W * src.channels;
  const size_t dstride = bgr ? h->bgr_stage_stride : h->pyr_stride;
  uint8_t* pyr0 = h->pyr2[set] + (size_t)img0 * h->pyr_stride;
  uint8_t* dst = bgr ? h->bgr_stage + (size_t)img0 * dstride : pyr0;
  const uint8_t* conv_src = dst;
  size_t conv_stride = dstride;
  int conv_pitch = dp;
  if (src.d_images) {
    const uint8_t* from = src.d_images + (size_t)img0 * src.img_stride;
    if (!bgr && h->lvl0[set] != nullptr) {
      // the caller's buffer IS level 0 (enqueue_batch checked pitch and alignment): nothing to copy
    } else if (bgr && W % 4 == 0 && src.pitch % 4 == 0 && src.img_stride % 4 == 0 && ((uintptr_t)src.d_images & 3) == 0) {
      conv_src = from; conv_stride = src.img_stride; conv_pitch = (int)src.pitch;   // convert in place, no staging copy
    } else if (src.pitch == (size_t)dp) {
      RPE_CK(cudaMemcpy2DAsync(dst, dstride, from, src.img_stride, src.pitch * H, ni, cudaMemcpyDeviceToDevice, s));
    } else {
      for (int i = 0; i < ni; ++i)
        RPE_CK(cudaMemcpy2DAsync(dst + (size_t)i * dstride, dp, from + (size_t)i * src.img_stride, src.pitch, row, H,
                                 cudaMemcpyDeviceToDevice, s));
    }
  } else if (src.host_contiguous) {
    RPE_CK(cudaMemcpy2DAsync(dst, dstride, src.host_image(0) + (ptrdiff_t)img0 * src.host_stride,
                             (size_t)src.host_stride, src.pitch * H, ni, cudaMemcpyHostToDevice, s));
  } else {
    for (int i = 0; i < ni; ++i)
      RPE_CK(cudaMemcpy2DAsync(dst + (size_t)i * dstride, dp, src.host_image(img0 + i), src.pitch, row, H,
                               cudaMemcpyHostToDevice, s));
  }
  if (bgr)
    launch_bgr2gray(conv_src, conv_stride, conv_pitch, pyr0, h->pyr_stride, p0, W, H, ni, h->gray_mode, s,
                    &h->launches[RPE_T_H2D]);
  return RPE_OK;
}

static void fill_host_source(BatchSource& src, const uint8_t* const* img1, const uint8_t* const* img2, int n_img, int H,
                             int W, size_t pitch, bool sequence, int channels) {
  src.img1 = img1; src.img2 = img2; src.pitch = pitch; src.sequence = sequence; src.channels = channels;
  // fast path: all images equally spaced in one host buffer with the device pitch -> one copy per chunk
  const int p0 = align_up(W * channels, 16);
  bool contiguous = (pitch == (size_t)p0) && n_img >= 2;
  const ptrdiff_t stride = contiguous ? src.host_image(1) - src.host_image(0) : 0;
  if (contiguous && stride >= (ptrdiff_t)(pitch * H)) {
    for (int i = 0; i < n_img && contiguous; ++i) contiguous = (src.host_image(i) == src.host_image(0) + (ptrdiff_t)i * stride);
  } else contiguous = false;
  src.host_contiguous = contiguous;
  src.host_stride = stride;
}

// The whole hot path over B pairs (pair batch: 2B images; sequence: B + 1 frames), enqueued without
// waiting.  Profiling: one chunk, every stage on the handle's stream with an event between stages.
// Otherwise the three-stream pipeline described at rpe_handle: uploads on s_copy, dense kernels of all
// chunks in order on the handle's stream, geometry tail + D2H on s_tail.
static int enqueue_batch(rpe_handle* h, const BatchSource& src, int B, int H, int W, const double K[9], double* R,
                         double* t, double* euler, int32_t* status, int32_t* info) {
  if (h->n_pending >= rpe_handle::kDepth) return fail(RPE_ERR_INVALID, "two submitted batches are pending: call rpe_wait_batch first");
  const bool prof = h->profiling;
  if (prof && h->n_pending) return fail(RPE_ERR_INVALID, "profiling needs one batch at a time: call rpe_wait_batch first");
  const Geometry* G;
  RPE_TRY(get_geometry(h, W, H, &G));
  const int pstep = src.sequence ? 1 : 2;
  const int n_img = src.sequence ? B + 1 : 2 * B;
  if (src.pitch < (size_t)W * src.channels) return fail(RPE_ERR_INVALID, "pitch smaller than one image row");
  if (src.d_images && src.img_stride < src.pitch * (size_t)H) return fail(RPE_ERR_INVALID, "img_stride smaller than one image");
  if (src.channels == 3) RPE_TRY(ensure_bgr_stage(h, h->NI, H, W));
  reset_profile(h);
  // buffer set of this batch; the batch that used it last (seq - 2) has been waited for (n_pending < 2), so its
  // pyramid slabs and tail buffers are free and the upload can start at once
  const int set = (int)(h->seq & 1);
  // device-resident gray frames in the pyramid's own layout (row pitch = align16(W), 16-byte aligned base and image
  // stride) are used in place as level 0: no D2D copy (0.48 ms of a 256-pair 720p step).  The frames must stay
  // valid and unchanged until rpe_wait_batch returns, as the header already requires.
  const bool alias = src.d_images && src.channels == 1 && src.pitch == (size_t)G->g.lv[0].pitch &&
                     ((uintptr_t)src.d_images & 15) == 0 && (src.img_stride & 15) == 0;
  h->lvl0[set] = alias ? src.d_images : nullptr;
  h->lvl0_stride[set] = alias ? src.img_stride : 0;
  cudaStream_t s_main = h->stream;
  cudaStream_t s_copy = (prof || src.d_images) ? s_main : h->s_copy;   // device inputs: a cheap D2D copy, in order
  cudaStream_t s_tail = prof ? s_main : h->s_tail;
  mark(h, RPE_T_H2D, prof);
  RPE_TRY(upload_K(h, set, K));
  RPE_TRY(clear_counters(h, s_main));
  const int chunk_img = (prof || h->n_workers < 2) ? n_img : std::min(n_img, 2 * h->chunk_pairs);
  const int n_chunks = (n_img + chunk_img - 1) / chunk_img;
  for (int c = 0; c < n_chunks; ++c) {
    const int img0 = c * chunk_img, ni = std::min(chunk_img, n_img - img0);
    // pairs completed by this chunk: [pair0, pair1); in a sequence the pair that straddles two chunks
    // belongs to the later one (same stream: its first frame's features are already there)
    const int pair0 = src.sequence ? std::max(img0 - 1, 0) : img0 / 2;
    const int pair1 = src.sequence ? img0 + ni - 1 : (img0 + ni) / 2;
    cudaEvent_t e_up = h->ev_up[c % rpe_handle::kChunkEvents];
    cudaEvent_t e_dense = h->ev_dense[c % rpe_handle::kChunkEvents];
    RPE_TRY(upload_images(h, *G, src, set, img0, ni, H, W, s_copy));
    if (s_copy != s_main) {
      RPE_CK(cudaEventRecord(e_up, s_copy));
      RPE_CK(cudaStreamWaitEvent(s_main, e_up, 0));
    }
    RPE_TRY(run_orb(h, *G, set, img0, ni, s_main, prof));
    if (pair1 > pair0) {
      RPE_TRY(run_match(h, set, pair0, pair1 - pair0, pstep, s_main, prof));
      if (s_tail != s_main) {
        RPE_CK(cudaEventRecord(e_dense, s_main));
        RPE_CK(cudaStreamWaitEvent(s_tail, e_dense, 0));
      }
      RPE_TRY(run_geometry(h, set, pair0, pair1 - pair0, s_tail, prof));
      mark(h, RPE_T_D2H, prof && c == n_chunks - 1);
      RPE_TRY(download_async(h, set, pair0, pair1 - pair0, s_tail));
    }
  }
  if (prof) cudaEventRecord(h->ev[RPE_T_COUNT], s_main);
  RPE_CK(cudaEventRecord(h->ev_done_call[set], s_tail));
  h->last_geom = G;
  h->last_pairs = B;
  h->last_pstep = pstep;
  rpe_handle::Pending& p = h->pend[h->n_pending++];
  p.B = B; p.set = set; p.R = R; p.t = t; p.euler = euler; p.status = status; p.info = info;
  ++h->seq;
  return RPE_OK;
}

// oldest pending batch
static int wait_batch(rpe_handle* h) {
  if (!h->n_pending) return fail(RPE_ERR_INVALID, "no submitted batch to wait for");
  const rpe_handle::Pending p = h->pend[0];
  h->pend[0] = h->pend[1];
  --h->n_pending;
  RPE_CK(cudaSetDevice(h->device));
  return finish_results(h, p);
}

// blocking form: nothing else may be pending, so that the results waited for are this call's
static int run_batch(rpe_handle* h, const BatchSource& src, int B, int H, int W, const double K[9], double* R,
                     double* t, double* euler, int32_t* status, int32_t* info) {
  if (h->n_pending) return fail(RPE_ERR_INVALID, "a submitted batch is pending: call rpe_wait_batch first");
  RPE_TRY(enqueue_batch(h, src, B, H, W, K, R, t, euler, status, info));
  return wait_batch(h);
}

extern "C" {

int rpe_estimate_batch(rpe_handle* h, const uint8_t* const* img1, const uint8_t* const* img2, int32_t B, int32_t H,
                       int32_t W, size_t pitch, const double K[9], double* R, double* t, double* euler,
                       int32_t* status, int32_t* info) {
  RPE_TRY(check_batch(h, B, H, W, pitch));
  if (!img1 || !img2 || !K) return fail(RPE_ERR_INVALID, "null argument");
  BatchSource src;
  fill_host_source(src, img1, img2, 2 * B, H, W, pitch, false, h->in_channels);
  return run_batch(h, src, B, H, W, K, R, t, euler, status, info);
}

int rpe_submit_batch(rpe_handle* h, const uint8_t* const* img1, const uint8_t* const* img2, int32_t B, int32_t H,
                     int32_t W, size_t pitch, const double K[9], double* R, double* t, double* euler,
                     int32_t* status, int32_t* info) {
  RPE_TRY(check_batch(h, B, H, W, pitch));
  if (!img1 || !img2 || !K) return fail(RPE_ERR_INVALID, "null argument");
  BatchSource src;
  fill_host_source(src, img1, img2, 2 * B, H, W, pitch, false, h->in_channels);
  return enqueue_batch(h, src, B, H, W, K, R, t, euler, status, info);
}

int rpe_submit_batch_device(rpe_handle* h, const uint8_t* d_images, size_t img_stride, int32_t B, int32_t H,
                            int32_t W, size_t pitch, const double K[9], double* R, double* t, double* euler,
                            int32_t* status, int32_t* info) {
  RPE_TRY(check_batch(h, B, H, W, pitch));
  if (!d_images || !K || img_stride < pitch * (size_t)H) return fail(RPE_ERR_INVALID, "bad device image arguments");
  BatchSource src;
  src.d_images = d_images; src.img_stride = img_stride; src.pitch = pitch; src.channels = h->in_channels;
  return enqueue_batch(h, src, B, H, W, K, R, t, euler, status, info);
}

int rpe_wait_batch(rpe_handle* h) {
  if (!h) return fail(RPE_ERR_INVALID, "null handle");
  return wait_batch(h);
}

int rpe_estimate_batch_device(rpe_handle* h, const uint8_t* d_images, size_t img_stride, int32_t B, int32_t H,
                              int32_t W, size_t pitch, const double K[9], double* R, double* t, double* euler,
                              int32_t* status, int32_t* info) {
  RPE_TRY(check_batch(h, B, H, W, pitch));
  if (!d_images || !K || img_stride < pitch * (size_t)H) return fail(RPE_ERR_INVALID, "bad device image arguments");
  BatchSource src;
  src.d_images = d_images; src.img_stride = img_stride; src.pitch = pitch; src.channels = h->in_channels;
  return run_batch(h, src, B, H, W, K, R, t, euler, status, info);
}

// ---- consecutive-frame form ----------------------------------------------------
static int check_sequence(rpe_handle* h, int F, int H, int W, size_t pitch) {
  if (!h) return fail(RPE_ERR_INVALID, "null handle");
  if (F < 2) return fail(RPE_ERR_INVALID, "a sequence needs at least 2 frames");
  if (F > h->NI) return fail(RPE_ERR_CAPACITY, "more frames than 2 * max_pairs image slots");
  return check_batch(h, F - 1, H, W, pitch);
}

int rpe_submit_sequence(rpe_handle* h, const uint8_t* const* frames, int32_t F, int32_t H, int32_t W, size_t pitch,
                        const double K[9], double* R, double* t, double* euler, int32_t* status, int32_t* info) {
  RPE_TRY(check_sequence(h, F, H, W, pitch));
  if (!frames || !K) return fail(RPE_ERR_INVALID, "null argument");
  BatchSource src;
  fill_host_source(src, frames, nullptr, F, H, W, pitch, true, h->in_channels);
  return enqueue_batch(h, src, F - 1, H, W, K, R, t, euler, status, info);
}

int rpe_estimate_sequence(rpe_handle* h, const uint8_t* const* frames, int32_t F, int32_t H, int32_t W, size_t pitch,
                          const double K[9], double* R, double* t, double* euler, int32_t* status, int32_t* info) {
  // blocking form: nothing else may be pending, so that the results waited for are this call's (as run_batch)
  if (h && h->n_pending) return fail(RPE_ERR_INVALID, "a submitted batch is pending: call rpe_wait_batch first");
  RPE_TRY(rpe_submit_sequence(h, frames, F, H, W, pitch, K, R, t, euler, status, info));
  return wait_batch(h);
}

int rpe_submit_sequence_device(rpe_handle* h, const uint8_t* d_frames, size_t img_stride, int32_t F, int32_t H,
                               int32_t W, size_t pitch, const double K[9], double* R, double* t, double* euler,
                               int32_t* status, int32_t* info) {
  RPE_TRY(check_sequence(h, F, H, W, pitch));
  if (!d_frames || !K || img_stride < pitch * (size_t)H) return fail(RPE_ERR_INVALID, "bad device image arguments");
  BatchSource src;
  src.d_images = d_frames; src.img_stride = img_stride; src.pitch = pitch; src.sequence = true; src.channels = h->in_channels;
  return enqueue_batch(h, src, F - 1, H, W, K, R, t, euler, status, info);
}

int rpe_estimate_sequence_device(rpe_handle* h, const uint8_t* d_frames, size_t img_stride, int32_t F, int32_t H,
                                 int32_t W, size_t pitch, const double K[9], double* R, double* t, double* euler,
                                 int32_t* status, int32_t* info) {
  if (h && h->n_pending) return fail(RPE_ERR_INVALID, "a submitted batch is pending: call rpe_wait_batch first");
  RPE_TRY(rpe_submit_sequence_device(h, d_frames, img_stride, F, H, W, pitch, K, R, t, euler, status, info));
  return wait_batch(h);
}

int rpe_orb_detect(rpe_handle* h, const uint8_t* img, int32_t H, int32_t W, size_t pitch, float* x, float* y,
                   float* size, float* angle, float* response, int32_t* octave, uint8_t* desc, int32_t* n_out,
                   int32_t* flags_out) {
  if (h && h->n_pending) return fail(RPE_ERR_INVALID, "a submitted batch is pending: call rpe_wait_batch first");
  RPE_TRY(check_batch(h, 1, H, W, pitch));
  if (!img || !n_out || pitch < (size_t)W) return fail(RPE_ERR_INVALID, "null argument / bad pitch");
  const Geometry* G;
  RPE_TRY(get_geometry(h, W, H, &G));
  reset_profile(h);
  h->lvl0[0] = nullptr;
  RPE_CK(cudaMemcpy2DAsync(h->pyr2[0], G->g.lv[0].pitch, img, pitch, W, H, cudaMemcpyHostToDevice, h->stream));
  RPE_TRY(clear_counters(h, h->stream));
  RPE_TRY(run_orb(h, *G, 0, 0, 1, h->stream, false));
  h->last_set = 0;
  h->last_geom = G;
  h->last_pairs = 0;
  h->last_pstep = 2;
  int32_t n = 0;
  uint32_t fl = 0;
  RPE_CK(cudaMemcpyAsync(&n, h->ob.kp_n, sizeof n, cudaMemcpyDeviceToHost, h->stream));
  RPE_CK(cudaMemcpyAsync(&fl, h->ob.img_flags, sizeof fl, cudaMemcpyDeviceToHost, h->stream));
  RPE_CK(cudaStreamSynchronize(h->stream));
  *n_out = n;
  if (flags_out) *flags_out = (int32_t)fl;
  auto get = [&](void* dst, const void* src, size_t bytes) -> cudaError_t {
    return dst ? cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost) : cudaSuccess;
  };
  RPE_CK(get(x, h->ob.kp_x, n * sizeof(float)));
  RPE_CK(get(y, h->ob.kp_y, n * sizeof(float)));
  RPE_CK(get(size, h->ob.kp_size, n * sizeof(float)));
  RPE_CK(get(angle, h->ob.kp_angle, n * sizeof(float)));
  RPE_CK(get(response, h->ob.kp_resp, n * sizeof(float)));
  RPE_CK(get(octave, h->ob.kp_oct, n * sizeof(int32_t)));
  RPE_CK(get(desc, h->ob.desc, (size_t)n * 32));
  return RPE_OK;
}

int rpe_bgr_to_gray(rpe_handle* h, const uint8_t* bgr, int32_t H, int32_t W, size_t pitch, int32_t gray_mode,
                    uint8_t* gray) {
  if (h && h->n_pending) return fail(RPE_ERR_INVALID, "a submitted batch is pending: call rpe_wait_batch first");
  RPE_TRY(check_batch(h, 1, H, W, pitch));
  if (!bgr || !gray || pitch < (size_t)3 * W || (gray_mode != RPE_GRAY_CVTCOLOR && gray_mode != RPE_GRAY_IMREAD_PNG))
    return fail(RPE_ERR_INVALID, "bad argument");
  const Geometry* G;
  RPE_TRY(get_geometry(h, W, H, &G));
  RPE_TRY(ensure_bgr_stage(h, 1, H, W));
  reset_profile(h);
  const int dp = align_up(3 * W, 16), p0 = G->g.lv[0].pitch;
  RPE_CK(cudaMemcpy2DAsync(h->bgr_stage, dp, bgr, pitch, (size_t)3 * W, H, cudaMemcpyHostToDevice, h->stream));
  launch_bgr2gray(h->bgr_stage, h->bgr_stage_stride, dp, h->pyr2[0], h->pyr_stride, p0, W, H, 1, gray_mode, h->stream,
                  &h->launches[RPE_T_H2D]);
  RPE_CK(cudaGetLastError());
  RPE_CK(cudaMemcpy2DAsync(gray, W, h->pyr2[0], p0, W, H, cudaMemcpyDeviceToHost, h->stream));
  RPE_CK(cudaStreamSynchronize(h->stream));
  return RPE_OK;
}

int rpe_match(rpe_handle* h, const uint8_t* desc1, int32_t n1, const uint8_t* desc2, int32_t n2, int32_t* matches,
              int32_t* m_out) {
  if (h && h->n_pending) return fail(RPE_ERR_INVALID, "a submitted batch is pending: call rpe_wait_batch first");
  if (!h || !m_out || n1 < 0 || n2 < 0) return fail(RPE_ERR_INVALID, "bad argument");
  if (n1 > h->kp_stride || n2 > h->kp_stride) return fail(RPE_ERR_CAPACITY, "more descriptors than the handle's keypoint capacity");
  RPE_CK(cudaSetDevice(h->device));
  reset_profile(h);
  cudaStream_t s = h->stream;
  const size_t KS = h->kp_stride;
  if (n1) RPE_CK(cudaMemcpyAsync(h->ob.desc, desc1, (size_t)n1 * 32, cudaMemcpyHostToDevice, s));
  if (n2) RPE_CK(cudaMemcpyAsync(h->ob.desc + KS * 32, desc2, (size_t)n2 * 32, cudaMemcpyHostToDevice, s));
  const int32_t nn[2] = {n1, n2};
  RPE_CK(cudaMemcpyAsync(h->ob.kp_n, nn, sizeof nn, cudaMemcpyHostToDevice, s));
  RPE_CK(cudaStreamSynchronize(s));   // nn is a stack buffer
  int* ln = h->launches;
  launch_match(h->ob, h->mb[0], h->kp_stride, h->cfg.nfeatures, 1, 2, h->cfg.match_mode, s, &ln[RPE_T_MATCH]);
  launch_build_matches(h->ob, h->mb[0], h->kp_stride, h->mcap, h->sort_cap_match, 1, 2, h->cfg.match_mode, h->cfg.ratio,
                       h->cfg.max_matches, s, &ln[RPE_T_MATCH_SORT]);
  RPE_CK(cudaGetLastError());
  int32_t m = 0;
  RPE_CK(cudaMemcpyAsync(&m, h->mb[0].m_n, sizeof m, cudaMemcpyDeviceToHost, s));
  RPE_CK(cudaStreamSynchronize(s));
  *m_out = m;
  if (matches && m) {
    std::vector<int32_t> q(m), t(m), d(m);
    RPE_CK(cudaMemcpy(q.data(), h->mb[0].m_q, m * sizeof(int32_t), cudaMemcpyDeviceToHost));
    RPE_CK(cudaMemcpy(t.data(), h->mb[0].m_t, m * sizeof(int32_t), cudaMemcpyDeviceToHost));
    RPE_CK(cudaMemcpy(d.data(), h->mb[0].m_d, m * sizeof(int32_t), cudaMemcpyDeviceToHost));
    for (int i = 0; i < m; ++i) { matches[3 * i] = q[i]; matches[3 * i + 1] = t[i]; matches[3 * i + 2] = d[i]; }
  }
  return RPE_OK;
}

static int upload_points(rpe_handle* h, const float* pts1, const float* pts2, int32_t m, const double K[9]) {
  if (m < 0 || m > h->mcap) return fail(RPE_ERR_CAPACITY, "more points than the handle's match capacity");
  cudaStream_t s = h->stream;
  if (m) {
    RPE_CK(cudaMemcpyAsync(h->mb[0].pts1, pts1, (size_t)m * 2 * sizeof(float), cudaMemcpyHostToDevice, s));
    RPE_CK(cudaMemcpyAsync(h->mb[0].pts2, pts2, (size_t)m * 2 * sizeof(float), cudaMemcpyHostToDevice, s));
  }
  RPE_CK(cudaMemcpyAsync(h->mb[0].m_n, &m, sizeof m, cudaMemcpyHostToDevice, s));
  RPE_TRY(upload_K(h, 0, K));
  RPE_CK(cudaStreamSynchronize(s));
  h->last_set = 0;
  int dummy = 0;
  launch_normalize(h->mb[0], h->gb[0], h->mcap, 1, h->dK[0], s, &dummy);
  return RPE_OK;
}

int rpe_find_essential(rpe_handle* h, const float* pts1, const float* pts2, int32_t m, const double K[9], double* E,
                       uint8_t* mask, int32_t* info) {
  if (h && h->n_pending) return fail(RPE_ERR_INVALID, "a submitted batch is pending: call rpe_wait_batch first");
  if (!h || !K || (m > 0 && (!pts1 || !pts2))) return fail(RPE_ERR_INVALID, "bad argument");
  RPE_CK(cudaSetDevice(h->device));
  reset_profile(h);
  if (!h->gb[0].tr_subset) {
    const int cap = (h->cfg.ransac_max_iters + 63) / 64 * 64;
    RPE_CK(cudaMalloc((void**)&h->gb[0].tr_subset, (size_t)cap * 5 * sizeof(int32_t)));
    RPE_CK(cudaMalloc((void**)&h->gb[0].tr_nmodels, (size_t)cap * sizeof(int32_t)));
    RPE_CK(cudaMalloc((void**)&h->gb[0].tr_counts, (size_t)cap * 10 * sizeof(int32_t)));
    RPE_CK(cudaMalloc((void**)&h->gb[0].tr_models, (size_t)cap * 90 * sizeof(double)));
    h->gb[0].tr_cap = cap;
  }
  RPE_TRY(upload_points(h, pts1, pts2, m, K));
  launch_ransac(h->mb[0], h->gb[0], h->mcap, 1, h->dK[0], ransac_params(h), h->stream, &h->launches[RPE_T_RANSAC]);
  RPE_CK(cudaGetLastError());
  RPE_CK(cudaStreamSynchronize(h->stream));
  if (E) RPE_CK(cudaMemcpy(E, h->gb[0].E, 9 * sizeof(double), cudaMemcpyDeviceToHost));
  if (mask && m) RPE_CK(cudaMemcpy(mask, h->gb[0].rmask, m, cudaMemcpyDeviceToHost));
  if (info) RPE_CK(cudaMemcpy(info, h->gb[0].rinfo, 4 * sizeof(int32_t), cudaMemcpyDeviceToHost));
  return RPE_OK;
}

int rpe_ransac_trace(rpe_handle* h, int32_t max_hyp, int32_t* subset, int32_t* nmodels, int32_t* counts,
                     double* models, int32_t* n_hyp) {
  if (!h || !h->gb[0].tr_subset || !n_hyp) return fail(RPE_ERR_INVALID, "no trace available");
  RPE_CK(cudaSetDevice(h->device));
  int32_t ri[4];
  RPE_CK(cudaMemcpy(ri, h->gb[0].rinfo, sizeof ri, cudaMemcpyDeviceToHost));
  const int n = std::min(std::min(ri[3], max_hyp), h->gb[0].tr_cap);
  *n_hyp = n;
  if (n <= 0) return RPE_OK;
  if (subset) RPE_CK(cudaMemcpy(subset, h->gb[0].tr_subset, (size_t)n * 5 * sizeof(int32_t), cudaMemcpyDeviceToHost));
  if (nmodels) RPE_CK(cudaMemcpy(nmodels, h->gb[0].tr_nmodels, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost));
  if (counts) RPE_CK(cudaMemcpy(counts, h->gb[0].tr_counts, (size_t)n * 10 * sizeof(int32_t), cudaMemcpyDeviceToHost));
  if (models) RPE_CK(cudaMemcpy(models, h->gb[0].tr_models, (size_t)n * 90 * sizeof(double), cudaMemcpyDeviceToHost));
  return RPE_OK;
}

int rpe_recover_pose(rpe_handle* h, const double E[9], const float* pts1, const float* pts2, int32_t m,
                     const double K[9], double* R, double* t, double* euler, int32_t* count, uint8_t* mask) {
  if (h && h->n_pending) return fail(RPE_ERR_INVALID, "a submitted batch is pending: call rpe_wait_batch first");
  if (!h || !E || !K || (m > 0 && (!pts1 || !pts2))) return fail(RPE_ERR_INVALID, "bad argument");
  RPE_CK(cudaSetDevice(h->device));
  reset_profile(h);
  RPE_CK(cudaMemcpy(h->gb[0].E, E, 9 * sizeof(double), cudaMemcpyHostToDevice));
  RPE_TRY(upload_points(h, pts1, pts2, m, K));
  GeomBuffers gb = h->gb[0];
  gb.rinfo = nullptr;
  launch_pose(h->mb[0], gb, h->mcap, 1, ransac_params(h), nullptr, h->stream, &h->launches[RPE_T_POSE]);
  RPE_CK(cudaGetLastError());
  RPE_CK(cudaStreamSynchronize(h->stream));
  if (R) RPE_CK(cudaMemcpy(R, h->gb[0].R, 9 * sizeof(double), cudaMemcpyDeviceToHost));
  if (t) RPE_CK(cudaMemcpy(t, h->gb[0].t, 3 * sizeof(double), cudaMemcpyDeviceToHost));
  if (euler) RPE_CK(cudaMemcpy(euler, h->gb[0].euler, 3 * sizeof(double), cudaMemcpyDeviceToHost));
  if (count) RPE_CK(cudaMemcpy(count, h->gb[0].pose_cnt, sizeof(int32_t), cudaMemcpyDeviceToHost));
  if (mask && m) RPE_CK(cudaMemcpy(mask, h->gb[0].pmask, m, cudaMemcpyDeviceToHost));
  return RPE_OK;
}

int rpe_estimate_from_matches(rpe_handle* h, const float* pts1, const float* pts2, const int32_t* counts, int32_t B,
                              int32_t cap, const double K[9], double* R, double* t, double* euler, int32_t* status,
                              int32_t* info) {
  if (h && h->n_pending) return fail(RPE_ERR_INVALID, "a submitted batch is pending: call rpe_wait_batch first");
  if (!h || !pts1 || !pts2 || !counts || !K || B < 1 || cap < 1) return fail(RPE_ERR_INVALID, "bad argument");
  if (B > h->NP) return fail(RPE_ERR_CAPACITY, "batch larger than max_pairs");
  if (cap > h->mcap) return fail(RPE_ERR_CAPACITY, "more points per pair than the handle's match capacity");
  for (int i = 0; i < B; ++i)
    if (counts[i] < 0 || counts[i] > cap) return fail(RPE_ERR_INVALID, "counts[i] outside [0, cap]");
  RPE_CK(cudaSetDevice(h->device));
  reset_profile(h);
  const bool prof = h->profiling;
  const int set = 0;
  cudaStream_t s = h->stream;
  const MatchBuffers& mb = h->mb[set];
  const size_t row = (size_t)cap * 2 * sizeof(float), dpitch = (size_t)h->mcap * 2 * sizeof(float);
  RPE_CK(cudaMemcpy2DAsync(mb.pts1, dpitch, pts1, row, row, B, cudaMemcpyHostToDevice, s));
  RPE_CK(cudaMemcpy2DAsync(mb.pts2, dpitch, pts2, row, row, B, cudaMemcpyHostToDevice, s));
  RPE_CK(cudaMemcpyAsync(mb.m_n, counts, (size_t)B * sizeof(int32_t), cudaMemcpyHostToDevice, s));
  RPE_CK(cudaMemsetAsync(mb.pstat, 0, (size_t)B * sizeof(int32_t), s));
  RPE_TRY(upload_K(h, set, K));
  RPE_TRY(run_geometry(h, set, 0, B, s, prof));
  if (prof) cudaEventRecord(h->ev[RPE_T_D2H], s);
  RPE_TRY(download_async(h, set, 0, B, s));
  RPE_CK(cudaEventRecord(h->ev_done_call[set], s));
  h->last_pairs = B;
  h->last_pstep = 2;
  rpe_handle::Pending p;
  p.B = B; p.set = set; p.R = R; p.t = t; p.euler = euler; p.status = status; p.info = info;
  const bool keep_prof = h->profiling;
  h->profiling = false;                       // only the two geometry stages carry events here
  const int rc = finish_results(h, p);
  h->profiling = keep_prof;
  if (rc != RPE_OK) return rc;
  if (prof) {
    RPE_CK(cudaEventElapsedTime(&h->ms[RPE_T_RANSAC], h->ev[RPE_T_RANSAC], h->ev[RPE_T_POSE]));
    RPE_CK(cudaEventElapsedTime(&h->ms[RPE_T_POSE], h->ev[RPE_T_POSE], h->ev[RPE_T_D2H]));
  }
  return RPE_OK;
}

int rpe_evaluate_sequence(rpe_handle* h, int32_t n, const double* est_R, const double* est_t, const double* est_euler,
                          const double* gt_euler, const double* gt_pos, int32_t euler_convention, double* errors,
                          double* summary) {
  if (h && h->n_pending) return fail(RPE_ERR_INVALID, "a submitted batch is pending: call rpe_wait_batch first");
  if (!h || n < 1 || !est_R || !est_t || !est_euler || !gt_euler || !gt_pos) return fail(RPE_ERR_INVALID, "bad argument");
  if (euler_convention != RPE_EULER_YUP && euler_convention != RPE_EULER_ZYX) return fail(RPE_ERR_INVALID, "bad euler convention");
  RPE_CK(cudaSetDevice(h->device));
  // scratch: inputs 9 + 3 + 3 + 3 + 3 doubles per frame, errors 5 per frame, summary 25; grown on demand
  const size_t per = 21 + 5, need = (size_t)n * per + 25;
  if (need > h->eval_cap) {
    if (h->eval_buf) cudaFree(h->eval_buf);
    h->eval_buf = nullptr; h->eval_cap = 0;
    RPE_CK(cudaMalloc((void**)&h->eval_buf, need * sizeof(double)));
    h->eval_cap = need;
  }
  double* d = h->eval_buf;
  double *dR = d, *dt = dR + 9 * (size_t)n, *de = dt + 3 * (size_t)n, *dg = de + 3 * (size_t)n, *dp = dg + 3 * (size_t)n,
         *derr = dp + 3 * (size_t)n, *dsum = derr + 5 * (size_t)n;
  cudaStream_t s = h->stream;
  RPE_CK(cudaMemcpyAsync(dR, est_R, 9 * (size_t)n * sizeof(double), cudaMemcpyHostToDevice, s));
  RPE_CK(cudaMemcpyAsync(dt, est_t, 3 * (size_t)n * sizeof(double), cudaMemcpyHostToDevice, s));
  RPE_CK(cudaMemcpyAsync(de, est_euler, 3 * (size_t)n * sizeof(double), cudaMemcpyHostToDevice, s));
  RPE_CK(cudaMemcpyAsync(dg, gt_euler, 3 * (size_t)n * sizeof(double), cudaMemcpyHostToDevice, s));
  RPE_CK(cudaMemcpyAsync(dp, gt_pos, 3 * (size_t)n * sizeof(double), cudaMemcpyHostToDevice, s));
  launch_evaluate(n, dR, dt, de, dg, dp, euler_convention, derr, dsum, s);
  RPE_CK(cudaGetLastError());
  if (errors) RPE_CK(cudaMemcpyAsync(errors, derr, 5 * (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, s));
  if (summary) RPE_CK(cudaMemcpyAsync(summary, dsum, 25 * sizeof(double), cudaMemcpyDeviceToHost, s));
  RPE_CK(cudaStreamSynchronize(s));
  return RPE_OK;
}

int rpe_summarize_errors(rpe_handle* h, int32_t n, const double* errors, double* summary) {
  if (h && h->n_pending) return fail(RPE_ERR_INVALID, "a submitted batch is pending: call rpe_wait_batch first");
  if (!h || n < 1 || !errors || !summary) return fail(RPE_ERR_INVALID, "bad argument");
  RPE_CK(cudaSetDevice(h->device));
  const size_t need = (size_t)n * 5 + 25;
  if (need > h->eval_cap) {
    if (h->eval_buf) cudaFree(h->eval_buf);
    h->eval_buf = nullptr; h->eval_cap = 0;
    RPE_CK(cudaMalloc((void**)&h->eval_buf, need * sizeof(double)));
    h->eval_cap = need;
  }
  double *derr = h->eval_buf, *dsum = derr + 5 * (size_t)n;
  RPE_CK(cudaMemcpyAsync(derr, errors, 5 * (size_t)n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  launch_eval_summary(n, derr, dsum, h->stream);
  RPE_CK(cudaGetLastError());
  RPE_CK(cudaMemcpyAsync(summary, dsum, 25 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  RPE_CK(cudaStreamSynchronize(h->stream));
  return RPE_OK;
}

int rpe_get_stage(rpe_handle* h, int32_t pair, int32_t view, int32_t stage, int32_t arg, void* out, size_t* n) {
  if (h && h->n_pending) return fail(RPE_ERR_INVALID, "a submitted batch is pending: call rpe_wait_batch first");
  if (!h || !n || !out) return fail(RPE_ERR_INVALID, "null argument");
  if (pair < 0 || pair >= h->NP || view < 0 || view > 1) return fail(RPE_ERR_INVALID, "bad pair/view");
  RPE_CK(cudaSetDevice(h->device));
  RPE_CK(cudaStreamSynchronize(h->stream));
  const size_t img = (size_t)h->last_pstep * pair + view, KS = h->kp_stride, MC = h->mcap;
  const size_t cap = *n;
  auto copy = [&](const void* src, size_t bytes) -> int {
    const size_t c = std::min(bytes, cap);
    if (c) RPE_CK(cudaMemcpy(out, src, c, cudaMemcpyDeviceToHost));
    *n = c;
    return RPE_OK;
  };
  int32_t kn = 0, mn = 0;
  RPE_CK(cudaMemcpy(&kn, h->ob.kp_n + img, sizeof kn, cudaMemcpyDeviceToHost));
  RPE_CK(cudaMemcpy(&mn, h->mb[h->last_set].m_n + pair, sizeof mn, cudaMemcpyDeviceToHost));
  switch (stage) {
    case RPE_STAGE_PYRAMID_LEVEL:
    case RPE_STAGE_BLURRED_LEVEL: {
      if (!h->last_geom || arg < 0 || arg >= h->cfg.nlevels) return fail(RPE_ERR_INVALID, "no geometry / bad level");
      const LevelGeom& L = h->last_geom->g.lv[arg];
      const uint8_t* base = (stage == RPE_STAGE_PYRAMID_LEVEL ? h->pyr2[h->last_set] : h->ob.blur) + img * h->pyr_stride + L.off;
      if (stage == RPE_STAGE_PYRAMID_LEVEL && arg == 0 && h->lvl0[h->last_set])
        base = h->lvl0[h->last_set] + img * h->lvl0_stride[h->last_set];
      const size_t bytes = (size_t)L.w * L.h;
      if (cap < bytes) return fail(RPE_ERR_INVALID, "buffer too small");
      if (stage == RPE_STAGE_BLURRED_LEVEL) {
        // blurred levels are stored in 16-column strips (rpe_device.cuh: blur_strip_offset): one strided copy per strip
        for (int x = 0; x < L.w; x += kBlurStripW)
          RPE_CK(cudaMemcpy2D((uint8_t*)out + x, L.w, base + blur_strip_offset(x, 0, L.h), kBlurStripW,
                              std::min(kBlurStripW, L.w - x), L.h, cudaMemcpyDeviceToHost));
      } else {
        RPE_CK(cudaMemcpy2D(out, L.w, base, L.pitch, L.w, L.h, cudaMemcpyDeviceToHost));
      }
      *n = bytes;
      return RPE_OK;
    }
    case RPE_STAGE_KP_X: return copy(h->ob.kp_x + img * KS, kn * sizeof(float));
    case RPE_STAGE_KP_Y: return copy(h->ob.kp_y + img * KS, kn * sizeof(float));
    case RPE_STAGE_KP_SIZE: return copy(h->ob.kp_size + img * KS, kn * sizeof(float));
    case RPE_STAGE_KP_ANGLE: return copy(h->ob.kp_angle + img * KS, kn * sizeof(float));
    case RPE_STAGE_KP_RESPONSE: return copy(h->ob.kp_resp + img * KS, kn * sizeof(float));
    case RPE_STAGE_KP_OCTAVE: return copy(h->ob.kp_oct + img * KS, kn * sizeof(int32_t));
    case RPE_STAGE_DESCRIPTORS: return copy(h->ob.desc + img * KS * 32, (size_t)kn * 32);
    case RPE_STAGE_MATCHES: {
      std::vector<int32_t> q(mn), t(mn), d(mn), o((size_t)mn * 3);
      if (mn) {
        RPE_CK(cudaMemcpy(q.data(), h->mb[h->last_set].m_q + pair * MC, mn * sizeof(int32_t), cudaMemcpyDeviceToHost));
        RPE_CK(cudaMemcpy(t.data(), h->mb[h->last_set].m_t + pair * MC, mn * sizeof(int32_t), cudaMemcpyDeviceToHost));
        RPE_CK(cudaMemcpy(d.data(), h->mb[h->last_set].m_d + pair * MC, mn * sizeof(int32_t), cudaMemcpyDeviceToHost));
      }
      for (int i = 0; i < mn; ++i) { o[3 * i] = q[i]; o[3 * i + 1] = t[i]; o[3 * i + 2] = d[i]; }
      const size_t c = std::min(o.size() * sizeof(int32_t), cap);
      memcpy(out, o.data(), c);
      *n = c;
      return RPE_OK;
    }
    case RPE_STAGE_PTS1: return copy(h->mb[h->last_set].pts1 + pair * MC * 2, (size_t)mn * 2 * sizeof(float));
    case RPE_STAGE_PTS2: return copy(h->mb[h->last_set].pts2 + pair * MC * 2, (size_t)mn * 2 * sizeof(float));
    case RPE_STAGE_E: return copy(h->gb[h->last_set].E + pair * 9, 9 * sizeof(double));
    case RPE_STAGE_RANSAC_MASK: return copy(h->gb[h->last_set].rmask + pair * MC, mn);
    case RPE_STAGE_POSE_MASK: return copy(h->gb[h->last_set].pmask + pair * MC, mn);
    default: return fail(RPE_ERR_INVALID, "unknown stage");
  }
}

}  // extern "C"
