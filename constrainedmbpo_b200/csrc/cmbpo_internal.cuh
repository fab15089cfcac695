// cmbpo_internal.cuh -- structures shared by the kernels and the C-ABI layer.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace cmbpo {

constexpr int kMaxLayers = 8;
constexpr int kMaxSeg = 3;     // input row = concat of up to 3 segments (obs | act | next_obs)
constexpr int kMaxMembers = 16;
constexpr int kTileRows = 128;  // rows per CTA tile (= TMEM lanes)
constexpr int kRegionCols = 256;  // one 256-column TMEM accumulator region per row-tile stream (2 streams)
constexpr int kMaxSlots = 8;      // weight ring slots

enum Activation : int { ACT_NONE = 0, ACT_SWISH = 1, ACT_RELU = 2, ACT_TANH = 3, ACT_SIGMOID = 4 };

// One FC layer of the ensemble (mbpo/models/fc.py:83-114) as the kernel sees it.
// B operand (= W^T, K-major) of each member is split in two N-halves, one per CTA of the pair.
struct LayerPlan {
  int k_real, n_real;   // in / out width of the layer
  int kpad, npad;       // kpad = roundup(k_real + 2, 16): two extra K rows carry the bias (hi+lo bf16)
  int act;              // Activation
  float prescale;       // 0.5 for swish/sigmoid (tanh form works on x/2), else 1
  uint32_t smem_off;    // (unused by the streaming kernel; layers go through the weight ring)
  uint32_t img_off;     // byte offset inside one (member, half) block of the packed image
  uint32_t bytes;       // (npad/2) * kpad * 2
};

struct ModelPlan {
  int E;                // ensemble members (bnn.py:96 num_nets)
  int L;                // FC layers including the head
  int in_dim;           // input width
  int n_head;           // width of the last layer (2*D if probabilistic else D)
  int D;                // output dim per head
  int probabilistic;    // head split into mean | logvar (bnn.py:111-112, :190-191)
  LayerPlan layer[kMaxLayers];
  uint32_t w_smem_bytes;  // w_slots * slot_bytes
  uint32_t blk_bytes;     // bytes of one (member, half) image block
  uint32_t a0_bytes;      // 128 * k0pad * 2: staged input tile of one stream
  uint32_t a_bytes;       // 128 * max_l>=1(kpad_l) * 2: activation tile of one stream (0 if L == 1)
  uint32_t slot_bytes;    // max_l(bytes_l) rounded up to 128: one weight ring slot
  uint32_t w_slots;       // ring depth (2..kMaxSlots), as many as fit beside the activation tiles
  uint32_t smem_total;    // dynamic smem the forward kernel needs
  uint32_t acc_cols;      // TMEM columns of one stream's hidden-layer accumulator = max hidden npad (streams at 0, acc_cols)
  uint32_t head_col;      // first TMEM column of the head accumulator shared by both streams (= 2 * acc_cols)
  uint32_t combo_off;     // byte offset of the NEXT member's layer-0 weights inside a "head + next layer 0" ring slot
};

// Head scratch layout ("tile-major"): the rows are cut in blocks of kTileRows (one CTA tile); block b
// holds [slot e][head column j][128 rows] contiguously, so a CTA's output for one (tile, member) is one
// contiguous n_head*512-byte region and consumers still read 128 consecutive rows per column.
__host__ __device__ inline size_t head_index(long long g, int e, int col, int E, int n_head) {
  return (((size_t)(g >> 7) * E + e) * n_head + col) * (size_t)kTileRows + (size_t)(g & (kTileRows - 1));
}

// Arguments of the input staging pass (K0): concat of the row segments -> input scaler -> bf16 ->
// canonical K-major A tiles of layer 0, one a0_bytes block per (input slot, 128-row block).
struct StageParams {
  const float* seg_ptr[kMaxSeg];       // input segments; row-major [slots?][B][width]
  int seg_width[kMaxSeg];
  long long seg_slot_stride[kMaxSeg];  // elements between input slots (3-D input), else 0
  int n_slots;                         // 1 (2-D input shared by all members) or E (3-D, fc.py:107-108)
  int in_dim, kpad;
  int chunked;                         // 1: wide-path image (K chunks of 64, no bias-carrier columns)
  int tf32;                            // 1 (with chunked): fp32 elements for kind::tf32, K chunks of 32 (same 16 KB blocks)
  int split3;                          // 1 (with tf32): columns [x_hi | x_lo | x_hi] against weight rows [W_hi ; W_hi ; W_lo]
  const float* in_mu;                  // [in_dim] scaler_in mean          (utils.py:161)
  const float* in_inv;                 // [in_dim] 1 / max(sqrt(var)*sc, 1e-2)
  const int* n_rows_dev;               // optional device scalar with the live row count
  long long n_rows;
  long long slot_stride_bytes;         // bytes between input slots of the image
  uint8_t* a0img;                      // out
};

// Arguments of one ensemble forward launch.
struct FwdParams {
  ModelPlan plan;
  const uint8_t* wimg;          // packed bf16 weight image [E][2][blk_bytes]
  const uint8_t* a0img;         // staged layer-0 A tiles [input slot][128-row block][a0_bytes]
  long long a0_slot_stride;     // bytes between input slots of a0img
  const float* head_scale;      // [n_head] s_out for mean cols, 1 for logvar cols (utils.py:172,192)
  const float* head_shift;      // [n_head] mu_out for mean cols, 2*log(s_out) for logvar cols
  int per_slot_input;           // 0: 2-D input shared by all members; 1: 3-D input (fc.py:107-108)
  int member_of_slot[kMaxMembers];  // weights used for output slot j
  int input_of_slot[kMaxMembers];   // input slot feeding output slot j (3-D only)
  const int* perm_dev;          // optional device [E] permutation overriding both maps (member = perm[j], input = perm[perm[j]])
  const int* n_rows_dev;        // optional device scalar with the live row count (rollout loop)
  long long n_rows;             // row count if n_rows_dev == nullptr
  float* heads;                 // out: tile-major head scratch (head_index), scaled mean | shifted logvar
  int* status;                  // watchdog / error word
  long long* trace;             // debug timeline, only read by -DCMBPO_TRACE builds
};

// ---------------------------------------------------------------------------------------
// Wide path (any layer width up to kWideMaxWidth, e.g. the reference's shipped (512, 512) nets,
// examples/config/antsafe/cmbpo_uc.py:36): one launch per FC layer, activations as bf16 tile images
// in HBM/L2 between the launches.  Image layout ("chunked canonical"): rows in blocks of 128, K in
// chunks of kWideKC columns; one (row block, K chunk) is a contiguous 16 KB block holding 8x8 core
// matrices: element (r, k) at (r/8)*1024 + (k/8)*128 + (r%8)*16 + (k%8)*2 -- the SWIZZLE_NONE K-major
// operand layout of tcgen05.mma with LBO = 128, SBO = 1024, so a stage is filled by one bulk copy.
// ---------------------------------------------------------------------------------------
constexpr int kWideKC = 64;            // K columns per pipeline stage
constexpr int kWideNC = 256;           // N columns per accumulator (one tcgen05.mma N)
constexpr int kWideMaxWidth = 2048;
constexpr uint32_t kWideBlk = 128u * kWideKC * 2u;   // 16 KB: one (128 rows, 64 k) block

struct WideLayerPlan {
  int k_real, n_real;
  int kc;                // K columns per 16 KB operand block: 64 (bf16) or 32 (tf32: fp32 elements, kind::tf32)
  int split3;            // tf32 layer 0: K axis = [x_hi | x_lo | x_hi] . [W_hi ; W_hi ; W_lo] (3xTF32 split, kpad = roundup(3 k_real, kc))
  int kpad;              // roundup(k_real, kc)
  int npad;              // hidden: roundup(n_real, kc) (= next layer's kpad); head: roundup(n_real, 16)
  int k_chunks, n_chunks;
  int act;
  float prescale;
  size_t w_off;          // byte offset of this layer in the wide weight image
  size_t b_off;          // float offset of this layer in the bias table ([E][npad], prescaled)
};

struct WideParams {
  WideLayerPlan lp;
  int E, n_head, is_head;
  int tf32;                    // operands are fp32 images consumed by tcgen05.mma kind::tf32
  const uint8_t* w_img;        // this layer: [slot e][n chunk][cta half][k chunk] blocks of kWideBlk bytes
  const float* bias;           // this layer: [E][npad]
  const uint8_t* a_img;        // input image [input slot][128-row block][k chunk]
  long long a_slot_stride;     // bytes between input slots (0: one input shared by all members)
  uint8_t* out_img;            // hidden: next layer's input image [slot e][128-row block][k chunk]
  long long out_slot_stride;
  float* heads;                // head: tile-major fp32 scratch (head_index)
  long long head_blk0;         // first 128-row block of this slab inside the head scratch
  const float* head_scale;
  const float* head_shift;
  int member_of_slot[kMaxMembers];
  int input_of_slot[kMaxMembers];
  const int* perm_dev;         // optional device [E] permutation: member = perm[j]; layer 0 (perm_input) input = perm[perm[j]]
  int perm_input;
  int per_slot_input;
  const int* n_rows_dev;       // optional live row count of the WHOLE batch
  long long n_rows;            // rows of the whole batch if n_rows_dev == nullptr
  long long row0;              // first row of this slab
  long long slab_rows;         // rows of this slab (multiple of 256 except for the last slab)
  int* status;
};

// Fused wide path (fused_wide.cu): the whole FC stack of a wide ensemble in one launch, activations kept in one
// shared-memory tile per CTA (a_blocks 16 KB blocks), weights streamed from the wide image through `stages` ring slots.
constexpr int kFusedWideMaxStages = 8;
constexpr int kFusedWideMaxABlocks = 8;      // widest hidden layer: 8 x 64 bf16 columns (or 8 x 32 tf32 columns)
constexpr int kFusedWideBudgetBlocks = 14;   // 14 x 16 KB = 224 KB of the 227 KB a CTA may use
constexpr int kFusedWideMinStages = 3;

struct FusedWideParams {
  WideLayerPlan lp[kMaxLayers];  // the wide path's layer plans (w_off / b_off relative to w_img / bias)
  int L, E, n_head, tf32;
  int a_blocks;                  // 16 KB blocks of the activation tile = max over hidden layers of npad / kc
  int a0_blocks;                 // 16 KB blocks of the staged input tile = layer 0's k_chunks
  int stages;                    // weight ring depth
  int streams;                   // row tiles per instance: 2 when every layer is one N chunk and two activation tiles fit, else 1
  const uint8_t* w_img;          // wide weight image (all layers)
  const float* bias;             // wide bias table (all layers)
  const uint8_t* a0_img;         // K0's chunked image [input slot][128-row block][k chunk]
  long long a0_slot_stride;
  float* heads;                  // out: tile-major fp32 head scratch (head_index)
  const float* head_scale;
  const float* head_shift;
  int member_of_slot[kMaxMembers];
  int input_of_slot[kMaxMembers];
  const int* perm_dev;           // optional device [E] permutation: member = perm[j], input = perm[perm[j]]
  int per_slot_input;
  const int* n_rows_dev;
  long long n_rows;
  int* status;
  long long* trace;              // debug timeline, only read by -DCMBPO_TRACE builds
};

// Slot maps of the 3-D ('iv_gae') path: host arrays by default, a device permutation when the caller drew it on the GPU
// (cmbpo_draw_permutation) so that a captured rollout replays with a fresh forward_shuffle every time.
#ifdef __CUDACC__
__device__ __forceinline__ int perm_at(const int* perm, int j, int E) {
  const int m = __ldg(perm + j);
  return m < 0 ? 0 : (m >= E ? E - 1 : m);
}
__device__ __forceinline__ int fwd_member_of(const FwdParams& p, int j) {
  return p.perm_dev ? perm_at(p.perm_dev, j, p.plan.E) : p.member_of_slot[j];
}
__device__ __forceinline__ int fwd_input_of(const FwdParams& p, int j) {
  return p.perm_dev ? perm_at(p.perm_dev, perm_at(p.perm_dev, j, p.plan.E), p.plan.E) : p.input_of_slot[j];
}
__device__ __forceinline__ int wide_member_of(const WideParams& p, int j) {
  return p.perm_dev ? perm_at(p.perm_dev, j, p.E) : p.member_of_slot[j];
}
__device__ __forceinline__ int wide_input_of(const WideParams& p, int j) {
  return (p.perm_dev && p.perm_input) ? perm_at(p.perm_dev, perm_at(p.perm_dev, j, p.E), p.E) : p.input_of_slot[j];
}
#endif


}  // namespace cmbpo
