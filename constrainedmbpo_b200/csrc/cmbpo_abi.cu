// cmbpo_abi.cu -- extern "C" surface of libcmbpo_b200.so (declared in include/cmbpo_b200.h).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <utility>
#include <vector>
#include "../../include/cmbpo_b200.h"
#include "cmbpo_internal.cuh"
#include "step_post.cuh"

namespace cmbpo {
bool plan_fwd_smem(ModelPlan& pl);
cudaError_t pack_layer(const float* W, const float* b, uint8_t* img, int E, const LayerPlan& lp,
                       uint32_t blk_bytes, cudaStream_t stream);
cudaError_t launch_ensemble_fwd(const FwdParams& p, int hidden_act, int num_sms, cudaStream_t stream);
cudaError_t launch_stage_inputs(const StageParams& p, int num_sms, cudaStream_t stream);
cudaError_t pack_wide_layer(const float* W, const float* b, uint8_t* img, float* bias_tab, int E,
                            const WideLayerPlan& lp, int tf32, cudaStream_t stream);
cudaError_t launch_wide_layer(const WideParams& p, int act, int num_sms, cudaStream_t stream);
cudaError_t launch_fused_wide(const FusedWideParams& p, int act, int num_sms, cudaStream_t stream);
}  // namespace cmbpo

using namespace cmbpo;
static_assert(kMaxMembers == CMBPO_MAX_MEMBERS && kMaxSeg == CMBPO_MAX_SEGMENTS, "include/cmbpo_b200.h and cmbpo_internal.cuh disagree");

static thread_local std::string g_err;
static int fail(int code, const std::string& msg) { g_err = msg; return code; }
static int fail_cuda(cudaError_t e, const char* where) {
  g_err = std::string(where) + ": " + cudaGetErrorString(e);
  return CMBPO_ERR_CUDA;
}
#define CU(x, where) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return fail_cuda(e_, where); } while (0)

struct Handle {
  cmbpo_cfg cfg;
  ModelPlan plan;
  int device = 0;
  int num_sms = 0;
  uint8_t* wimg = nullptr;
  float* consts = nullptr;     // in_mu | in_inv | head_scale | head_shift
  int* elites_dev = nullptr;
  int n_elites = 0;
  int* status_dev = nullptr;   // device alias of status_host (mapped pinned memory: readable after a trap)
  volatile int* status_host = nullptr;
  unsigned int* ticket_dev = nullptr;   // step_post's last-block counter
  uint32_t layers_set = 0;
  int hidden_act = 0;
  bool scalers_set = false;
  bool profiling = false;
  // wide path (wide_fwd.cu): any layer wider than the fused kernel's 254 columns, or CMBPO_FORCE_WIDE=1
  bool wide = false;
  bool tf32 = false;           // cfg.operand_dtype == CMBPO_DTYPE_TF32: wide path with fp32 operand images
  WideLayerPlan wl[kMaxLayers];
  uint8_t* wimg_wide = nullptr;
  float* bias_wide = nullptr;
  int wide_max_kc = 0;         // widest hidden activation image, in 64-column chunks
  // fused wide path (fused_wide.cu): hidden layers up to 8 operand blocks wide run the whole stack in one launch
  bool fused_wide = false;
  int fw_a_blocks = 0, fw_a0_blocks = 0, fw_stages = 0, fw_streams = 1;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events;
};

static inline int roundup(int x, int m) { return (x + m - 1) / m * m; }
constexpr long long kWideSlabRows = 131072;   // rows per pass of the wide path (bounds the activation images)

extern "C" const char* cmbpo_last_error(void) { return g_err.c_str(); }

extern "C" int cmbpo_device_ok(int device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || device >= n) { cudaGetLastError(); return 0; }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { cudaGetLastError(); return 0; }
  return prop.major == 10 ? 1 : 0;
}

extern "C" int cmbpo_create(const cmbpo_cfg* cfg, int device, void** handle) {
  if (!cfg || !handle) return fail(CMBPO_ERR_ARG, "cmbpo_create: null argument");
  *handle = nullptr;
  const int E = cfg->num_members, L = cfg->num_layers;
  if (E < 1 || E > 8) return fail(CMBPO_ERR_UNSUPPORTED, "num_members must be in [1, 8]");
  if (L < 1 || L > kMaxLayers) return fail(CMBPO_ERR_UNSUPPORTED, "num_layers must be in [1, 8]");
  if (cfg->in_dim < 1 || cfg->out_dim < 1) return fail(CMBPO_ERR_ARG, "in_dim/out_dim must be positive");
  const int n_head = cfg->probabilistic ? 2 * cfg->out_dim : cfg->out_dim;
  if (cfg->layer_width[L - 1] != n_head)
    return fail(CMBPO_ERR_ARG, "last layer width must be out_dim (x2 if probabilistic), bnn.py:190-191");
  if (cfg->layer_act[L - 1] != CMBPO_ACT_NONE)
    return fail(CMBPO_ERR_UNSUPPORTED, "output activation must be None on this path (bnn.py:194-196)");
  if (!cmbpo_device_ok(device))
    return fail(CMBPO_ERR_CUDA, "no sm_100 CUDA device: this library has no CPU or other-arch fallback");

  Handle* h = new Handle();
  h->cfg = *cfg;
  h->device = device;
  ModelPlan& pl = h->plan;
  memset(&pl, 0, sizeof(pl));
  pl.E = E; pl.L = L; pl.in_dim = cfg->in_dim; pl.n_head = n_head; pl.D = cfg->out_dim;
  pl.probabilistic = cfg->probabilistic ? 1 : 0;
  uint32_t off = 0;
  for (int l = 0; l < L; ++l) {
    LayerPlan& lp = pl.layer[l];
    lp.k_real = (l == 0) ? cfg->in_dim : cfg->layer_width[l - 1];
    lp.n_real = cfg->layer_width[l];
    if (lp.n_real < 1) { delete h; return fail(CMBPO_ERR_ARG, "layer width must be positive"); }
    lp.kpad = roundup(lp.k_real + 2, 16);
    lp.npad = (l < L - 1) ? roundup(lp.n_real + 2, 16) : roundup(lp.n_real, 16);
    lp.act = cfg->layer_act[l];
    if (lp.act < 0 || lp.act > CMBPO_ACT_SIGMOID) { delete h; return fail(CMBPO_ERR_ARG, "unknown activation id"); }
    lp.prescale = (lp.act == CMBPO_ACT_SWISH || lp.act == CMBPO_ACT_SIGMOID) ? 0.5f : 1.0f;
    if (lp.kpad > kRegionCols || lp.npad > kRegionCols) h->wide = true;   // beyond one 256-column accumulator per stream
    if (lp.k_real > kWideMaxWidth || lp.n_real > kWideMaxWidth) {
      delete h;
      return fail(CMBPO_ERR_UNSUPPORTED, "layer wider than 2048 columns");
    }
    lp.bytes = (uint32_t)(lp.npad / 2) * lp.kpad * 2u;
    lp.smem_off = off;
    lp.img_off = off;
    off += lp.bytes;
  }
  h->hidden_act = (L > 1) ? cfg->layer_act[0] : CMBPO_ACT_NONE;
  for (int l = 1; l < L - 1; ++l)
    if (cfg->layer_act[l] != h->hidden_act) {
      delete h;
      return fail(CMBPO_ERR_UNSUPPORTED, "hidden layers must share one activation (constructor.py:57-60)");
    }
  pl.blk_bytes = off;
  if (cfg->operand_dtype != CMBPO_DTYPE_BF16 && cfg->operand_dtype != CMBPO_DTYPE_TF32) {
    delete h;
    return fail(CMBPO_ERR_ARG, "operand_dtype must be CMBPO_DTYPE_BF16 or CMBPO_DTYPE_TF32");
  }
  if (cfg->operand_dtype == CMBPO_DTYPE_TF32) { h->tf32 = true; h->wide = true; }
  if (const char* fw = getenv("CMBPO_FORCE_WIDE")) if (fw[0] == '1') h->wide = true;   // test hook: same nets on both kernels
  if (!h->wide && !plan_fwd_smem(pl)) h->wide = true;   // shapes beyond the fused kernel's shared-memory budget
  size_t wide_img = 0, wide_bias = 0;
  if (h->wide) {
    if (roundup(n_head, 16) > kWideNC) {
      delete h;
      return fail(CMBPO_ERR_UNSUPPORTED, "head wider than 256 columns");
    }
    for (int l = 0; l < L; ++l) {
      WideLayerPlan& w = h->wl[l];
      w.k_real = pl.layer[l].k_real; w.n_real = pl.layer[l].n_real;
      w.kc = h->tf32 ? kWideKC / 2 : kWideKC;
      // tf32 layer 0: the scaled inputs reach |x| ~ 1e2 where the input scaler hits its 1e-2 floor (utils.py:161), and a
      // 10-bit mantissa cannot hold them (nor their products with rounded weights) to 1e-3 of the output scale: the first
      // GEMM runs as x_hi W_hi + x_lo W_hi + x_hi W_lo (3xTF32), laid out as one GEMM over a K axis three times as long
      w.split3 = (h->tf32 && l == 0) ? 1 : 0;
      w.kpad = roundup((w.split3 ? 3 : 1) * w.k_real, w.kc);
      w.npad = (l < L - 1) ? roundup(w.n_real, w.kc) : roundup(w.n_real, 16);
      w.k_chunks = w.kpad / w.kc;
      w.n_chunks = (w.npad + kWideNC - 1) / kWideNC;
      w.act = pl.layer[l].act;
      w.prescale = h->tf32 ? 1.0f : pl.layer[l].prescale;   // tf32 path: exact sigmoid / tanh, no x/2 form
      w.w_off = wide_img; w.b_off = wide_bias;
      wide_img += (size_t)E * w.n_chunks * 2 * w.k_chunks * kWideBlk;
      wide_bias += (size_t)E * w.npad;
      if (l < L - 1 && w.npad / w.kc > h->wide_max_kc) h->wide_max_kc = w.npad / w.kc;
    }
    // whole-stack kernel when one activation tile (+ the staged input tile + a weight ring) fits in shared memory;
    // CMBPO_WIDE_FUSED=0 keeps the layer-at-a-time kernel (test hook: the two agree bit for bit)
    h->fw_a_blocks = h->wide_max_kc;
    h->fw_a0_blocks = h->wl[0].k_chunks;
    // two row tiles per instance when every layer is one N chunk (<= 256 columns) and both tiles' buffers fit
    bool one_chunk = true;
    for (int l = 0; l < L; ++l) if (h->wl[l].n_chunks != 1) one_chunk = false;
    h->fw_streams = (one_chunk && 2 * (h->fw_a_blocks + h->fw_a0_blocks) + kFusedWideMinStages <= kFusedWideBudgetBlocks) ? 2 : 1;
    if (const char* fs = getenv("CMBPO_FW_STREAMS")) if (fs[0] == '1') h->fw_streams = 1;   // measurement / test hook
    h->fw_stages = kFusedWideBudgetBlocks - h->fw_streams * (h->fw_a_blocks + h->fw_a0_blocks);
    if (h->fw_stages > kFusedWideMaxStages) h->fw_stages = kFusedWideMaxStages;
    if (const char* fs = getenv("CMBPO_FW_STAGES")) {   // measurement hook: a shallower weight ring
      const int v = atoi(fs);
      if (v >= 2 && v < h->fw_stages) h->fw_stages = v;
    }
    h->fused_wide = L >= 2 && h->fw_a_blocks >= 1 && h->fw_a_blocks <= kFusedWideMaxABlocks && h->fw_stages >= kFusedWideMinStages;
    if (const char* ff = getenv("CMBPO_WIDE_FUSED")) if (ff[0] == '0') h->fused_wide = false;
  }
  CU(cudaSetDevice(device), "cudaSetDevice");
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, device), "cudaGetDeviceProperties");
  h->num_sms = prop.multiProcessorCount;
  if (h->wide) {
    CU(cudaMalloc(&h->wimg_wide, wide_img), "cudaMalloc(wimg_wide)");
    CU(cudaMemset(h->wimg_wide, 0, wide_img), "cudaMemset(wimg_wide)");
    CU(cudaMalloc(&h->bias_wide, wide_bias * sizeof(float)), "cudaMalloc(bias_wide)");
    CU(cudaMemset(h->bias_wide, 0, wide_bias * sizeof(float)), "cudaMemset(bias_wide)");
  } else {
    CU(cudaMalloc(&h->wimg, (size_t)E * 2 * pl.blk_bytes), "cudaMalloc(wimg)");
    CU(cudaMemset(h->wimg, 0, (size_t)E * 2 * pl.blk_bytes), "cudaMemset(wimg)");
  }
  CU(cudaMalloc(&h->consts, sizeof(float) * (2 * pl.in_dim + 2 * n_head)), "cudaMalloc(consts)");
  CU(cudaMalloc(&h->elites_dev, sizeof(int) * 64), "cudaMalloc(elites)");
  // watchdog word in mapped host memory: the kernels write the wait code and __trap(), which poisons the context --
  // a device-memory word could not be copied back afterwards, the host mapping can still be read
  {
    int* hp = nullptr;
    CU(cudaHostAlloc(&hp, sizeof(int), cudaHostAllocMapped), "cudaHostAlloc(status)");
    *hp = 0;
    h->status_host = hp;
    CU(cudaHostGetDevicePointer(&h->status_dev, hp, 0), "cudaHostGetDevicePointer(status)");
  }
  CU(cudaMalloc(&h->ticket_dev, sizeof(unsigned int)), "cudaMalloc(ticket)");
  CU(cudaMemset(h->ticket_dev, 0, sizeof(unsigned int)), "cudaMemset(ticket)");
  // defaults: identity scalers (utils.py:108-115), every member elite
  int rc = cmbpo_set_scalers(h, nullptr, nullptr, nullptr, nullptr, 1.0f, nullptr);
  if (rc != 0) { cmbpo_destroy(h); return rc; }
  std::vector<int32_t> all(E);
  for (int e = 0; e < E; ++e) all[e] = e;
  rc = cmbpo_set_elites(h, all.data(), E, nullptr);
  if (rc != 0) { cmbpo_destroy(h); return rc; }
  h->scalers_set = false;
  *handle = h;
  return CMBPO_OK;
}

extern "C" void cmbpo_destroy(void* handle) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h) return;
  cudaSetDevice(h->device);
  cudaFree(h->wimg); cudaFree(h->wimg_wide); cudaFree(h->bias_wide); cudaFree(h->consts); cudaFree(h->elites_dev); cudaFreeHost((void*)h->status_host); cudaFree(h->ticket_dev);
  delete h;
}

extern "C" int cmbpo_set_weights(void* handle, int32_t layer, const float* w_dev, const float* b_dev, void* stream) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h || !w_dev) return fail(CMBPO_ERR_ARG, "cmbpo_set_weights: null argument");
  if (layer < 0 || layer >= h->plan.L) return fail(CMBPO_ERR_ARG, "cmbpo_set_weights: layer out of range");
  CU(cudaSetDevice(h->device), "cudaSetDevice");
  if (h->wide) {
    CU(pack_wide_layer(w_dev, b_dev, h->wimg_wide + h->wl[layer].w_off, h->bias_wide + h->wl[layer].b_off, h->plan.E,
                       h->wl[layer], h->tf32 ? 1 : 0, (cudaStream_t)stream), "pack_wide_layer");
  } else {
    CU(pack_layer(w_dev, b_dev, h->wimg, h->plan.E, h->plan.layer[layer], h->plan.blk_bytes, (cudaStream_t)stream),
       "pack_layer");
  }
  h->layers_set |= (1u << layer);
  return CMBPO_OK;
}

extern "C" int cmbpo_set_scalers(void* handle, const float* mu_in, const float* var_in, const float* mu_out,
                                 const float* var_out, float sc_factor, void* stream) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h) return fail(CMBPO_ERR_ARG, "cmbpo_set_scalers: null handle");
  if ((mu_in == nullptr) != (var_in == nullptr) || (mu_out == nullptr) != (var_out == nullptr))
    return fail(CMBPO_ERR_ARG, "cmbpo_set_scalers: mu and var must be given together");
  const ModelPlan& pl = h->plan;
  std::vector<float> c(2 * pl.in_dim + 2 * pl.n_head);
  float* in_mu = c.data();
  float* in_inv = in_mu + pl.in_dim;
  float* hs = in_inv + pl.in_dim;
  float* hb = hs + pl.n_head;
  for (int i = 0; i < pl.in_dim; ++i) {
    in_mu[i] = mu_in ? mu_in[i] : 0.f;
    // utils.py:161  max(sqrt(var)*sc_factor, 1e-2)
    const float s = var_in ? fmaxf(sqrtf(var_in[i]) * sc_factor, 1e-2f) : 1.f;
    in_inv[i] = 1.0f / s;
  }
  for (int j = 0; j < pl.n_head; ++j) {
    const int d = j % pl.D;
    const float s = var_out ? fmaxf(sqrtf(var_out[d]) * sc_factor, 1e-2f) : 1.f;
    if (j < pl.D) {            // mean columns: utils.py:172
      hs[j] = s; hb[j] = mu_out ? mu_out[d] : 0.f;
    } else {                   // log-variance columns: utils.py:192
      hs[j] = 1.f; hb[j] = var_out ? 2.0f * logf(s) : 0.f;
    }
  }
  CU(cudaSetDevice(h->device), "cudaSetDevice");
  CU(cudaMemcpyAsync(h->consts, c.data(), c.size() * sizeof(float), cudaMemcpyHostToDevice, (cudaStream_t)stream),
     "cudaMemcpyAsync(consts)");
  CU(cudaStreamSynchronize((cudaStream_t)stream), "cudaStreamSynchronize");
  h->scalers_set = true;
  return CMBPO_OK;
}

extern "C" int cmbpo_set_elites(void* handle, const int32_t* elites, int32_t k, void* stream) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h || !elites) return fail(CMBPO_ERR_ARG, "cmbpo_set_elites: null argument");
  if (k < 1 || k > 64) return fail(CMBPO_ERR_ARG, "cmbpo_set_elites: k must be in [1, 64]");
  for (int i = 0; i < k; ++i)
    if (elites[i] < 0 || elites[i] >= h->plan.E) return fail(CMBPO_ERR_ARG, "cmbpo_set_elites: index out of range");
  CU(cudaSetDevice(h->device), "cudaSetDevice");
  CU(cudaMemcpyAsync(h->elites_dev, elites, sizeof(int32_t) * k, cudaMemcpyHostToDevice, (cudaStream_t)stream),
     "cudaMemcpyAsync(elites)");
  CU(cudaStreamSynchronize((cudaStream_t)stream), "cudaStreamSynchronize");
  h->n_elites = k;
  return CMBPO_OK;
}

// rows of the tile-major head scratch (cmbpo_internal.cuh head_index): whole 128-row blocks
static inline long long row_stride_for(long long n_rows) { return (n_rows + kTileRows - 1) / kTileRows * kTileRows; }

// workspace = head scratch | step_post partial sums | staged layer-0 tiles (E input slots)
static size_t a0_slot_bytes(const Handle* h, long long rows) {
  const long long rows_pad = (rows + 2 * kTileRows - 1) / (2 * kTileRows) * (2 * kTileRows);
  return (size_t)rows_pad * (h->wide ? h->wl[0].kpad : h->plan.layer[0].kpad) * (h->tf32 ? 4u : 2u);
}
// wide path: two ping-pong activation images [E slots][slab 128-row blocks][widest hidden K chunks]
static long long wide_slab_pad(long long rows) {
  const long long r = rows < kWideSlabRows ? rows : kWideSlabRows;
  return (r + 2 * kTileRows - 1) / (2 * kTileRows) * (2 * kTileRows);
}
static size_t wide_act_bytes(const Handle* h, long long rows) {
  if (!h->wide || h->fused_wide) return 0;   // the fused wide kernel keeps the activations in shared memory
  return (size_t)h->plan.E * (size_t)(wide_slab_pad(rows) / kTileRows) * h->wide_max_kc * kWideBlk;
}
static size_t ws_partials_offset(const Handle* h, long long rows) {
  return (size_t)h->plan.E * h->plan.n_head * row_stride_for(rows) * sizeof(float);
}
static size_t ws_a0_offset(const Handle* h, long long rows) {
  const size_t o = ws_partials_offset(h, rows) + (size_t)step_post_blocks(rows, h->num_sms) * 4 * sizeof(float);
  return (o + 255) & ~(size_t)255;
}

extern "C" size_t cmbpo_workspace_bytes(void* handle, int64_t max_rows) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h || max_rows < 0) return 0;
  const ModelPlan& pl = h->plan;
  return ws_a0_offset(h, max_rows) + (size_t)pl.E * a0_slot_bytes(h, max_rows) + 256 + 2 * wide_act_bytes(h, max_rows) + 256;
}

static int check_ready(Handle* h) {
  if (h->layers_set != (1u << h->plan.L) - 1u)
    return fail(CMBPO_ERR_STATE, "weights of every layer must be set before predict/step");
  return CMBPO_OK;
}

static int run_forward(Handle* h, const float* const* seg, const int32_t* seg_width, int nseg, long long n_rows,
                       const int* n_rows_dev, int ndim, const int32_t* perm, const int32_t* perm_dev, float* heads,
                       cudaStream_t stream) {
  const ModelPlan& pl = h->plan;
  FwdParams p;
  memset(&p, 0, sizeof(p));
  p.plan = pl;
  p.wimg = h->wimg;
  p.head_scale = h->consts + 2 * pl.in_dim;
  p.head_shift = h->consts + 2 * pl.in_dim + pl.n_head;
  StageParams sp;
  memset(&sp, 0, sizeof(sp));
  int wsum = 0;
  for (int s = 0; s < kMaxSeg; ++s) {
    sp.seg_ptr[s] = (s < nseg) ? seg[s] : nullptr;
    sp.seg_width[s] = (s < nseg && seg[s]) ? seg_width[s] : 0;
    sp.seg_slot_stride[s] = (ndim == 3) ? n_rows * (long long)sp.seg_width[s] : 0;
    wsum += sp.seg_width[s];
  }
  if (wsum != pl.in_dim) return fail(CMBPO_ERR_ARG, "Invalid input dimension.");  // fc.py:111
  // compact the segment list (a null middle segment would break the concat walk of the staging kernel)
  for (int s = 0, d = 0; s < kMaxSeg; ++s)
    if (sp.seg_width[s] > 0) {
      if (d != s) {
        sp.seg_ptr[d] = sp.seg_ptr[s]; sp.seg_width[d] = sp.seg_width[s]; sp.seg_slot_stride[d] = sp.seg_slot_stride[s];
        sp.seg_ptr[s] = nullptr; sp.seg_width[s] = 0; sp.seg_slot_stride[s] = 0;
      }
      ++d;
    }
  sp.n_slots = (ndim == 3) ? pl.E : 1;
  sp.in_dim = pl.in_dim; sp.kpad = h->wide ? h->wl[0].kpad : pl.layer[0].kpad;
  sp.chunked = h->wide ? 1 : 0;
  sp.tf32 = h->tf32 ? 1 : 0;
  sp.split3 = h->tf32 ? 1 : 0;
  if (h->tf32 && sp.kpad > 384) return fail(CMBPO_ERR_UNSUPPORTED, "tf32 operands: input width above 128 columns");
  sp.in_mu = h->consts; sp.in_inv = h->consts + pl.in_dim;
  sp.n_rows_dev = n_rows_dev; sp.n_rows = n_rows;
  sp.slot_stride_bytes = (long long)a0_slot_bytes(h, n_rows);
  sp.a0img = reinterpret_cast<uint8_t*>(heads) + ws_a0_offset(h, n_rows);
  p.a0img = sp.a0img;
  p.a0_slot_stride = sp.slot_stride_bytes;
  p.per_slot_input = (ndim == 3) ? 1 : 0;
  for (int e = 0; e < pl.E; ++e) {
    const int m = (ndim == 3 && perm) ? perm[e] : e;
    if (m < 0 || m >= pl.E) return fail(CMBPO_ERR_ARG, "perm entry out of range");
    p.member_of_slot[e] = m;
    p.input_of_slot[e] = (ndim == 3 && perm) ? perm[m] : e;
  }
  p.perm_dev = (ndim == 3) ? perm_dev : nullptr;
  p.n_rows_dev = n_rows_dev;
  p.n_rows = n_rows;
  p.heads = heads;
  p.status = h->status_dev;
#ifdef CMBPO_TRACE
  static long long* trace_dev = nullptr;
  const char* trace_path = getenv("CMBPO_TRACE_OUT");   // debug timeline of cluster 0 (tools/trace_report.py)
  if (trace_path) {
    if (!trace_dev) CU(cudaMalloc(&trace_dev, 4 * 4096 * sizeof(long long)), "cudaMalloc(trace)");
    CU(cudaMemsetAsync(trace_dev, 0, 4 * 4096 * sizeof(long long), stream), "memset(trace)");
    p.trace = trace_dev;
  }
#endif
  CU(launch_stage_inputs(sp, h->num_sms, stream), "launch_stage_inputs");
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  if (h->profiling) {
    CU(cudaEventCreate(&ev0), "cudaEventCreate");
    CU(cudaEventCreate(&ev1), "cudaEventCreate");
    CU(cudaEventRecord(ev0, stream), "cudaEventRecord");
  }
  if (!h->wide) {
    CU(launch_ensemble_fwd(p, h->hidden_act, h->num_sms, stream), "launch_ensemble_fwd");
  } else if (h->fused_wide) {
    // fused wide path: the whole stack in one launch, activations in shared memory
    FusedWideParams fp;
    memset(&fp, 0, sizeof(fp));
    for (int l = 0; l < pl.L; ++l) fp.lp[l] = h->wl[l];
    fp.L = pl.L; fp.E = pl.E; fp.n_head = pl.n_head; fp.tf32 = h->tf32 ? 1 : 0;
    fp.a_blocks = h->fw_a_blocks; fp.a0_blocks = h->fw_a0_blocks; fp.stages = h->fw_stages; fp.streams = h->fw_streams;
    fp.w_img = h->wimg_wide; fp.bias = h->bias_wide;
    fp.a0_img = sp.a0img; fp.a0_slot_stride = sp.slot_stride_bytes;
    fp.heads = heads; fp.head_scale = p.head_scale; fp.head_shift = p.head_shift;
    for (int e = 0; e < pl.E; ++e) { fp.member_of_slot[e] = p.member_of_slot[e]; fp.input_of_slot[e] = p.input_of_slot[e]; }
    fp.perm_dev = p.perm_dev;
    fp.per_slot_input = p.per_slot_input;
    fp.n_rows_dev = n_rows_dev; fp.n_rows = n_rows;
    fp.status = h->status_dev;
    fp.trace = p.trace;
    CU(launch_fused_wide(fp, h->hidden_act, h->num_sms, stream), "launch_fused_wide");
  } else {
    // wide path: per row slab, one launch per FC layer; activations ping-pong between two bf16 images
    const size_t act_bytes = wide_act_bytes(h, n_rows);
    uint8_t* act0 = sp.a0img + (((size_t)pl.E * a0_slot_bytes(h, n_rows) + 255) & ~(size_t)255);
    uint8_t* actbuf[2] = {act0, act0 + ((act_bytes + 255) & ~(size_t)255)};
    const long long slab_blocks = wide_slab_pad(n_rows) / kTileRows;
    for (long long row0 = 0; row0 < n_rows; row0 += kWideSlabRows) {
      const long long slab_rows = (n_rows - row0 < kWideSlabRows) ? n_rows - row0 : kWideSlabRows;
      for (int l = 0; l < pl.L; ++l) {
        WideParams wp;
        memset(&wp, 0, sizeof(wp));
        wp.lp = h->wl[l];
        wp.E = pl.E; wp.n_head = pl.n_head; wp.is_head = (l == pl.L - 1) ? 1 : 0;
        wp.tf32 = h->tf32 ? 1 : 0;
        wp.w_img = h->wimg_wide + h->wl[l].w_off;
        wp.bias = h->bias_wide + h->wl[l].b_off;
        for (int e = 0; e < pl.E; ++e) {
          wp.member_of_slot[e] = p.member_of_slot[e];
          wp.input_of_slot[e] = (l == 0) ? p.input_of_slot[e] : e;
        }
        wp.perm_dev = p.perm_dev;
        wp.perm_input = (l == 0) ? 1 : 0;
        if (l == 0) {
          wp.a_img = sp.a0img + (size_t)(row0 / kTileRows) * h->wl[0].k_chunks * kWideBlk;
          wp.a_slot_stride = sp.slot_stride_bytes;
          wp.per_slot_input = p.per_slot_input;
        } else {
          wp.a_img = actbuf[(l - 1) & 1];
          wp.a_slot_stride = (long long)slab_blocks * h->wl[l].k_chunks * kWideBlk;
          wp.per_slot_input = 1;
        }
        wp.out_img = actbuf[l & 1];
        wp.out_slot_stride = (long long)slab_blocks * (h->wl[l].npad / h->wl[l].kc) * kWideBlk;
        wp.heads = heads;
        wp.head_blk0 = row0 / kTileRows;
        wp.head_scale = p.head_scale; wp.head_shift = p.head_shift;
        wp.n_rows_dev = n_rows_dev; wp.n_rows = n_rows;
        wp.row0 = row0; wp.slab_rows = slab_rows;
        wp.status = h->status_dev;
        CU(launch_wide_layer(wp, h->wl[l].act, h->num_sms, stream), "launch_wide_layer");
      }
    }
  }
  if (h->profiling) {
    CU(cudaEventRecord(ev1, stream), "cudaEventRecord");
    h->prof_events.emplace_back(ev0, ev1);
  }
#ifdef CMBPO_TRACE
  if (trace_path) {
    std::vector<long long> host(4 * 4096);
    CU(cudaStreamSynchronize(stream), "sync(trace)");
    CU(cudaMemcpy(host.data(), trace_dev, host.size() * sizeof(long long), cudaMemcpyDeviceToHost), "memcpy(trace)");
    if (FILE* f = fopen(trace_path, "wb")) { fwrite(host.data(), sizeof(long long), host.size(), f); fclose(f); }
  }
#endif
  return CMBPO_OK;
}

extern "C" int cmbpo_predict(void* handle, const cmbpo_predict_io* io, void* stream_) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h || !io) return fail(CMBPO_ERR_ARG, "cmbpo_predict: null argument");
  int rc = check_ready(h);
  if (rc) return rc;
  if (io->ndim != 2 && io->ndim != 3) return fail(CMBPO_ERR_ARG, "Invalid input dimension.");
  if (io->n_rows < 0) return fail(CMBPO_ERR_ARG, "cmbpo_predict: negative row count");
  if (io->n_rows == 0) return CMBPO_OK;   // empty batch: empty outputs, nothing to launch
  if (!io->mean_dev) return fail(CMBPO_ERR_ARG, "cmbpo_predict: null output pointer");
  if (io->workspace_bytes < cmbpo_workspace_bytes(h, io->n_rows) || !io->workspace_dev)
    return fail(CMBPO_ERR_WORKSPACE, "cmbpo_predict: workspace too small");
  if (io->n_rows == 0) return CMBPO_OK;
  cudaStream_t stream = (cudaStream_t)stream_;
  CU(cudaSetDevice(h->device), "cudaSetDevice");
  const ModelPlan& pl = h->plan;
  float* heads = static_cast<float*>(io->workspace_dev);
  rc = run_forward(h, io->seg_dev, io->seg_width, CMBPO_MAX_SEGMENTS, io->n_rows, io->n_rows_dev, io->ndim, nullptr, nullptr, heads, stream);
  if (rc) return rc;
  const long long stride = row_stride_for(io->n_rows);
  if (io->agg == 0 || io->ndim == 3) {
    CU(launch_finalize_predict(heads, stride, io->n_rows, pl.E, pl.D, pl.probabilistic, io->mean_dev, io->var_dev,
                               h->num_sms, stream), "finalize_predict");
  } else {
    CU(launch_aggregate_predict(heads, stride, io->n_rows, pl.E, pl.D, pl.probabilistic, io->mean_dev, io->var_dev,
                                h->num_sms, stream), "aggregate_predict");
  }
  return CMBPO_OK;
}

extern "C" int cmbpo_step(void* handle, const cmbpo_step_io* io, void* stream_) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h || !io) return fail(CMBPO_ERR_ARG, "cmbpo_step: null argument");
  int rc = check_ready(h);
  if (rc) return rc;
  const ModelPlan& pl = h->plan;
  if (!pl.probabilistic) return fail(CMBPO_ERR_STATE, "cmbpo_step needs a probabilistic dynamics ensemble (fake_env.py:59)");
  if (io->ndim != 2 && io->ndim != 3) return fail(CMBPO_ERR_ARG, "obs/act must be 2-D or 3-D (fake_env.py:132-139)");
  if (io->n_rows == 0) return CMBPO_OK;
  const int obs_dim = pl.D - 1, act_dim = pl.in_dim - obs_dim;
  if (act_dim < 0) return fail(CMBPO_ERR_ARG, "in_dim smaller than obs_dim");
  if (obs_dim > 64) return fail(CMBPO_ERR_UNSUPPORTED, "obs_dim > 64 not supported by the post-processing tile");
  if (!io->obs_dev || (act_dim > 0 && !io->act_dev) || !io->next_obs_dev || !io->rew_dev || !io->term_dev ||
      !io->dkl_path_dev || !io->scalars_dev)
    return fail(CMBPO_ERR_ARG, "cmbpo_step: null input/output pointer");
  if (io->n_rows < 0) return fail(CMBPO_ERR_ARG, "cmbpo_step: negative row count");
  if (io->workspace_bytes < cmbpo_workspace_bytes(h, io->n_rows) || !io->workspace_dev)
    return fail(CMBPO_ERR_WORKSPACE, "cmbpo_step: workspace too small");
  if (io->env_id < 0 || io->env_id > CMBPO_ENV_WALKER2D) return fail(CMBPO_ERR_ARG, "unknown env id");
  if (io->n_rows == 0) return CMBPO_OK;
  cudaStream_t stream = (cudaStream_t)stream_;
  CU(cudaSetDevice(h->device), "cudaSetDevice");
  float* heads = static_cast<float*>(io->workspace_dev);
  const long long stride = row_stride_for(io->n_rows);
  float* partials = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(heads) + ws_partials_offset(h, io->n_rows));
  const float* segs[CMBPO_MAX_SEGMENTS] = {io->obs_dev, io->act_dev, nullptr};
  const int32_t widths[CMBPO_MAX_SEGMENTS] = {obs_dim, act_dim, 0};
  rc = run_forward(h, segs, widths, 2, io->n_rows, io->n_rows_dev, io->ndim, io->perm, io->perm_dev, heads, stream);
  if (rc) return rc;
  StepParams sp;
  memset(&sp, 0, sizeof(sp));
  sp.heads = heads; sp.row_stride = stride; sp.D = pl.D; sp.obs_dim = obs_dim;
  sp.per_slot = io->ndim == 3;
  sp.obs = io->obs_dev;
  sp.elite_idx = io->elite_idx_dev; sp.row_ids = io->row_id_dev;
  sp.elites = h->elites_dev; sp.n_elites = h->n_elites;
  sp.seed = io->seed; sp.rng_offset = io->rng_offset; sp.row_offset = io->row_offset;
  sp.rng_offset_dev = reinterpret_cast<const unsigned long long*>(io->rng_offset_dev);
  sp.noise = io->noise_dev;
  sp.sample_mode = io->sample_mode; sp.bound_logvar = io->bound_logvar;
  sp.max_logvar = io->max_logvar; sp.min_logvar = io->min_logvar;
  sp.head_shift = h->consts + 2 * pl.in_dim + pl.n_head;
  sp.env_id = io->env_id; sp.cost_mode = io->cost_mode;
  sp.n_rows_dev = io->n_rows_dev; sp.n_rows = io->n_rows;
  sp.next_obs = io->next_obs_dev; sp.rew = io->rew_dev; sp.term = io->term_dev; sp.cost = io->cost_dev;
  sp.dkl_path = io->dkl_path_dev; sp.ep_var_mean = io->ep_var_mean_dev; sp.ep_var = io->ep_var_dev;
  sp.partials = partials;
  sp.ticket = h->ticket_dev;
  sp.scalars = io->scalars_dev;
  const int n_blocks = step_post_blocks(io->n_rows, h->num_sms);
  CU(launch_step_post(sp, pl.E, n_blocks, stream), "launch_step_post");

  return CMBPO_OK;
}

extern "C" int cmbpo_forward_path(void* handle) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h) return fail(CMBPO_ERR_ARG, "cmbpo_forward_path: null handle");
  return !h->wide ? 0 : (h->fused_wide ? 2 : 1);
}

extern "C" int cmbpo_profile(void* handle, int enable) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h) return fail(CMBPO_ERR_ARG, "cmbpo_profile: null handle");
  h->profiling = enable != 0;
  return CMBPO_OK;
}

extern "C" int cmbpo_profile_read(void* handle, double* fwd_ms_total, int64_t* fwd_launches) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h || !fwd_ms_total || !fwd_launches) return fail(CMBPO_ERR_ARG, "cmbpo_profile_read: null argument");
  double total = 0.0;
  for (auto& pr : h->prof_events) {
    CU(cudaEventSynchronize(pr.second), "cudaEventSynchronize");
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, pr.first, pr.second), "cudaEventElapsedTime");
    total += ms;
    cudaEventDestroy(pr.first); cudaEventDestroy(pr.second);
  }
  *fwd_ms_total = total;
  *fwd_launches = (int64_t)h->prof_events.size();
  h->prof_events.clear();
  return CMBPO_OK;
}

extern "C" int cmbpo_status(void* handle, void* stream) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h) return fail(CMBPO_ERR_ARG, "cmbpo_status: null handle");
  const cudaError_t e = cudaStreamSynchronize((cudaStream_t)stream);
  const int st = h->status_host ? *h->status_host : 0;   // mapped host word: valid even if the context is poisoned
  if (st != 0) {
    char buf[160];
    snprintf(buf, sizeof(buf), "device watchdog tripped: wait code %d%s%s", st & 0xFFFF,
             e != cudaSuccess ? "; stream: " : "", e != cudaSuccess ? cudaGetErrorString(e) : "");
    return fail(CMBPO_ERR_CUDA, buf);
  }
  if (e != cudaSuccess) return fail_cuda(e, "cmbpo_status: stream");
  return CMBPO_OK;
}
