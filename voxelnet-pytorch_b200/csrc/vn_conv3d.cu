// Conv3d(k=3) + BatchNorm3d(eval) + ReLU as a bf16 implicit GEMM on tcgen05 / TMEM, TMA-fed.
// Replaces ConvMD(3, ...).forward (reference voxelnet/model.py:158-167) for the three middle layers
// instantiated at model.py:207-209, and (out_mode 1) the reshape at model.py:262.
//
// GEMM view: M = output positions, N = 64 output channels, K = 27 taps x Cin.
//
// Units of work (DESIGN.md "Conv3d"):
//   tile         16 (h) x 8 (w) output positions of one frame, all output depths: up to Dout
//                accumulators of 128 x 64 fp32 in TMEM (64 columns each).
//   input box    what one (tile, input plane p, kw) needs: 64 ch x 8 w x 18 h (h-halo'd, shifted by
//                kw-1 in w) - ONE TMA box, 128-byte rows, SWIZZLE_128B.  The three kh taps read it at
//                +kh*1024 B (one image row of the tile = one 8-row swizzle atom).  H / W borders are
//                TMA out-of-bounds zero fill.
//   live box     a box that may hold anything other than the input's background value (all-zero for
//                the scattered voxel grid).  Dead boxes are never loaded or multiplied; what they
//                would have contributed is a per-channel constant folded into the epilogue's shift
//                table (see k_fold_background).  Without a flag table every box is live.
//   group        consecutive tiles of one tile row whose *touched* accumulators (those fed by at
//                least one live box) fit the 8 TMEM slots.  A group shares one pass over the weights:
//                for each (64-channel chunk, kw) the 72 KB of weights of that pass are made resident
//                in shared memory once and every live box of every tile in the group is multiplied
//                against them.  k_conv3d_plan builds the groups on the device; the persistent CTAs
//                pull them from an atomic counter (sparse scenes are badly balanced otherwise).
//
// An input plane p feeds output depth d' = (p + pad_d - kd) / stride_d for every valid kd; those d'
// are consecutive, so the depth taps of one plane are fused along N: ONE tcgen05.mma of shape
// 128 x (64*nkd) x 16 multiplies the A tile with the nkd weight tiles laid side by side and lands in
// nkd adjacent accumulators (A is read from shared memory once instead of nkd times; N = 64 alone is
// shared-memory-bandwidth bound).
//
//   B operand: weights pre-packed as [chunk][kw][kd class][kh][kd descending = d' ascending] K-major
//              [64 n x 64 k] tiles already in the swizzled shared-memory image.
//   Warp roles: warp 0 = scheduler + TMA producer, warp 1 = TMEM allocator + MMA issuer (both run
//              warp-uniform control flow and elect one lane for the asynchronous instructions),
//              warps 2..5 = epilogue (TMEM -> registers -> scale/shift/ReLU -> global).
#include <cuda.h>
#include <cuda_bf16.h>
#include <string.h>

#include "vn_common.cuh"

namespace vn {
namespace {

constexpr int kTileH = 16, kTileW = 8;             // 128 output positions per accumulator
constexpr int kHaloH = kTileH + 2;
constexpr uint32_t kABytes = kHaloH * kTileW * 128;  // 18432: 144 rows of 64 bf16
constexpr uint32_t kBTapBytes = 64 * 128;            // 8192: [64 n][64 k] bf16
constexpr uint32_t kBSlotBytes = 9 * kBTapBytes;     // all weights of one (chunk, kw): 3 kh x 3 kd tiles = 72 KB
constexpr int kNA = 4, kNB = 2, kSched = 4;          // ring depths: A boxes, weight passes, group ids
constexpr int kSlots = 8;                            // 8 x 64 = 512 TMEM columns
constexpr int kMaxPlanes = 10;                       // live mask = 3 bits per plane in a uint32
constexpr int kMaxGroupTiles = 8;
constexpr int kMaxTilesW = 128;
constexpr int kMaxTableRows = 27;                    // shift-table rows (variants x Dout) kept in smem
constexpr int kThreads = 10 * 32;            // producer, MMA issuer, 8 epilogue warps
constexpr int kFill = 4;                              // box-filling warps of k_conv3d_bn_relu<.., kGather>: message i -> warp i % kFill
constexpr int kThreadsGather = kThreads + kFill * 32;
constexpr int kGatherTableRows = 8;                   // kGather: shift-table rows kept in smem; the rest of Control::table holds the mailboxes
constexpr uint32_t kSmemData = kNA * kABytes + kNB * kBSlotBytes;  // 221184

// What one input plane contributes (host-computed; see plan_planes()).
struct PlaneInfo {
  int32_t nkd;        // number of valid depth taps (0: plane unused)
  int32_t dmin;       // first output depth fed; tap j feeds dmin + j
  uint32_t woff;      // byte offset of this plane's first weight tile inside a (chunk, kw) weight group
  uint32_t wstride;   // bytes between the kh = 0, 1, 2 tile rows of the plane's kd class
};

// Work plan in the caller's workspace: header | groups[n_tiles] | tiles[n_tiles] | gtiles[n_tiles][kMaxGroupTiles]
// (gtiles[g][i] = copy of the plan of the group's i-th tile: the conv kernel fetches a group with ONE dependent access).
struct PlanHeader {
  int32_t n_groups;    // written by k_conv3d_plan (zeroed before it)
  int32_t next_group;  // the conv kernel's work counter (zeroed before it)
  int32_t pad[2];
};
struct GroupDesc {
  int32_t tile0;   // first tile (global tile index = (b * tiles_h + th) * tiles_w + tw)
  int32_t ntiles;  // consecutive tiles of the same tile row
  int32_t b;       // frame ...
  int16_t th, tw0; // ... tile row and first tile column (= tile0 decomposed: no integer divisions in the conv kernel)
};
struct TilePlan {
  uint32_t live;    // bit p*3 + kw: input box (p, kw) of this tile is live
  uint8_t touched;  // bit d: accumulator d receives at least one live box
  uint8_t slot0;    // first TMEM slot of the tile inside its group; accumulator d sits at
                    // slot0 + popcount(touched & ((1 << d) - 1))
  uint8_t nslots;   // popcount(touched)
  uint8_t restore;  // bit d: output block d holds stale content of an earlier call and no accumulator feeds
                    // it now: rewrite it with the layer background (see PlanParams::out_state)
};
static_assert(sizeof(TilePlan) == 8 && sizeof(GroupDesc) == 16 && sizeof(PlanHeader) == 16, "plan layout");

struct ConvParams {
  int B, Din, Dout, H, W;
  int nchunk;      // Cin / 64
  int tiles_h, tiles_w, n_tiles;
  int out_mode;    // 0: bf16 NDHWC, 1: f32 [B, 64*Dout, H, W] with channel = c*Dout + d, 2: bf16 [B, H, W, 64*Dout] with channel = d*64 + c
  int nvar;        // 1: shift rows [Dout][64]; 9: [3 h-classes][3 w-classes][Dout][64]; 25: [5][5][Dout][64]
  int table_smem;  // every table row is in shared memory (else only the interior variant: border rows come from global memory)
  int sub_classes; // 1: sub [Dout][64]; 9: sub [3][3][Dout][64] indexed like the 9-variant shift table
  const uint8_t* wpack;
  const float* scale;   // [64]
  const float* table;   // [nvar][Dout][64] (or [64] when table_plain)
  int reverse;          // take the groups of the plan last to first (VN_CONV_FLAGS_REVERSE)
  int out_ctotal, out_coff;  // out_mode 2: channels per position of y and the first channel this launch writes
  int out_up, out_oy, out_ox;  // out_mode 2: output position (h, w) is written at (h * out_up + out_oy, w * out_up + out_ox) of a
                               // [B, H * out_up, W * out_up, out_ctotal] tensor (a transposed convolution with kernel = stride)
  float relu_floor;     // y = max(conv * scale + shift, relu_floor) - sub: 0 = ReLU, -inf = no activation
  int pointwise;        // kS2 instantiation: 1x1 convolution - only the centre tap (kw = 1, kh = 1) exists, stride 1
  int table_plain;      // table is a plain [64] shift vector used for every row
  const float* sub;     // [Dout][64] subtracted after the ReLU, or nullptr
  void* out;
  PlanHeader* hdr;
  const GroupDesc* groups;
  const TilePlan* tiles;
  const TilePlan* gtiles;
  PlaneInfo plane[kMaxPlanes];
  unsigned long long* prof;  // kDebug instantiation only: [grid][16] cycle counters (vn_conv3d_debug_profile)
  int debug;                 // kDebug instantiation only: VN_DBG_* experiment switches (results are then wrong)
  unsigned long long* trace; // kDebug: [trace_n][4] per message of CTA 0: producer acquire / issuer saw / issued / flags
  int trace_n;
  double* stats;             // training kernel: [2][64] per-channel sum / sum of squares of the raw output
  // kGather instantiation: the input is never a dense tensor.  Cell (b, p, h, w) holds row
  //   voxel_offsets[b] + prefix[b][word] + popc(bits below) of `rows` if its bit is set in the voxelizer's occupancy bitmap
  // (cell index (p*H + h)*W + w inside the frame; rows at or beyond min(voxel_offsets[B], max_voxels) do not exist), else zero.
  const uint32_t* g_bitmap;     // [B][g_words]
  const uint32_t* g_prefix;     // [B][g_words] exclusive popcount prefix inside the frame
  const int32_t* g_voxoff;      // [B+1]
  const __nv_bfloat16* g_rows;  // [max_voxels][128]
  long long g_max_voxels;
  uint32_t g_words;
};

struct PlanParams {
  int B, Din, Dout, H, W, tiles_h, tiles_w;
  uint8_t* box_flags;  // [B][Din][tiles_h][tiles_w][3] or nullptr (every box live)
  int only_kw1;        // 1x1 convolution: only the boxes of the centre column shift exist
  int consume;         // reset the flags that were read
  int force_ring;      // boxes containing an H/W border position are live (input deviates there)
  // [B][tiles_h][tiles_w] bytes kept by the caller next to the OUTPUT buffer, or nullptr.  Bit d set:
  // output block (tile, d) holds something other than what the layer writes there when no live box
  // reaches it (a constant that depends on the layer parameters only).  With a state table the kernel
  // writes only blocks that are fed now or were fed by the previous call; without, every block.
  uint8_t* out_state;
  PlanHeader* hdr;
  GroupDesc* groups;
  TilePlan* tiles;
  TilePlan* gtiles;
  PlaneInfo plane[kMaxPlanes];
};

struct Control {
  uint64_t full_a[kNA], empty_a[kNA];
  uint64_t full_b[kNB], empty_b[kNB];
  uint64_t acc_full[kSlots], acc_empty[kSlots];
  uint64_t sched_full[kSched], sched_empty[kSched];
  int32_t sched[kSched];
  GroupDesc sched_gd[kSched];                     // the scheduled group's descriptor ...
  TilePlan sched_tiles[kSched][kMaxGroupTiles];   // ... and tile plans, published by the producer for the epilogue
  uint32_t tmem_base;
  uint32_t pad[1];
  uint32_t row_zero[kSlots];  // [dout]: relu(interior shift row) - sub == 0 for all channels
  uint4 msg[kNA];  // one message per A stage, written by the producer before it arms full_a (see kMsg*)
  float scale[64];
  float sub[kSlots * 64];
  alignas(16) float table[kMaxTableRows * 64];  // training kernel: rows 16..19 hold the CTA-level partial sums
};
// kGather: message i (a box or a control message) goes through the one-slot mailbox of filling warp i % kFill, which
// fills stage i % kNA and owns its full / empty barriers from then on: {first column, first row, plane | chunk << 8 |
// kind << 16, frame} + the issuer message.  dirty[s][lane] bit k: box row lane + 32 k of stage s holds a voxel row.
// Lives in Control::table behind the kGatherTableRows rows the first layer's epilogue needs.
struct Mailboxes {
  uint64_t full[kFill], empty[kFill];
  int4 req[kFill];
  uint4 msg[kFill];
  uint8_t dirty[kNA][32];
  uint32_t turn[kNA];  // uses of the stage armed so far (two warps take turns on a stage: see the wait in the filling loop)
};
static_assert(kFill % kNA == 0, "filling warps per stage");
static_assert(sizeof(Mailboxes) <= (kMaxTableRows - kGatherTableRows) * 64 * sizeof(float), "mailboxes fit behind the table rows");
__device__ __forceinline__ Mailboxes* ctl_mail(Control* ctl) { return reinterpret_cast<Mailboxes*>(&ctl->table[kGatherTableRows * 64]); }
__device__ __forceinline__ double* ctl_stat(Control* ctl) { return reinterpret_cast<double*>(&ctl->table[16 * 64]); }
constexpr uint32_t kSmemBytes = kSmemData + 1024 /*align slack*/ + (uint32_t)sizeof(Control);
static_assert(kSmemBytes <= 232448, "shared memory budget (227 KB)");

// Producer -> MMA issuer messages (Control::msg[stage]).  The producer does all the bookkeeping
// (which accumulators a box feeds, whether they are fresh, when a tile is complete); the issuing warp
// only interprets: x = flags, y = TMEM address of the first accumulator, z / w = (offset of the plane's
// first weight tile, distance between its kh rows) >> 4.  Everything that can be decided ahead of time is decided by
// the producer (it has slack; the issuing thread's instructions between two MMAs are on the critical path).
constexpr uint32_t kMsgNkdMask = 3;      // bits 0-1: depth taps fused along N (0: control message, no A box)
constexpr uint32_t kMsgNewPass = 4;      // first box of a weight pass: release the previous slot, wait for the next
constexpr uint32_t kMsgFreshShift = 4;   // bits 4-6: tap j starts its accumulator (accumulate = 0 on the first k-step)
constexpr uint32_t kMsgEnd = 0x80;       // no more work
constexpr uint32_t kMsgSplit = 1u << 24;       // the taps of this box are in different states (some fresh, some not): rare
constexpr uint32_t kMsgAccumulate = 1u << 25;  // no tap is fresh: the first MMA accumulates
constexpr uint32_t kMsgCommitShift = 8;  // bits 8-15: TMEM slots that are complete after this message
constexpr uint32_t kMsgWaitShift = 16;   // bits 16-23: TMEM slots the epilogue must have drained before the MMAs
constexpr uint32_t kMsgKhSkipShift = 26;  // bits 26-27 (kS2): how many of the two extra kh rows of the box this message does NOT have
constexpr uint32_t kABytesS2 = 17 * kTileW * 128;  // a stride-2 box: 17 sampled rows x 8 sampled columns x 64 channels
// experiment switches of the kDebug instantiation (tools/prof_conv.py): what the kernel costs WITHOUT ...
constexpr int kDbgNoALoad = 1;   // ... the TMA loads of the input boxes (the MMAs read stale shared memory)
constexpr int kDbgNoBLoad = 2;   // ... the bulk loads of the weights
constexpr int kDbgNoMma = 4;     // ... the tcgen05.mma instructions (commits only)
constexpr int kDbgNoStore = 8;   // ... the epilogue's global stores
constexpr int kDbgNoFence = 16;  // kGather: ... the filling warps' proxy fence
constexpr int kDbgNoLookup = 32; // kGather: ... the occupancy lookups (every box is treated as empty)
constexpr int kReqBox = 0, kReqControl = 1, kReqStop = 2;  // kGather mailbox kinds (Control::req[].z >> 16)
constexpr uint32_t kDescHi = 0x40004040u;  // SW128 K-major descriptor, high word: SBO = 1024 B, version 1, layout 2

__device__ __forceinline__ uint64_t desc_from_lo(uint32_t lo) { return ((uint64_t)kDescHi << 32) | lo; }

__host__ __device__ __forceinline__ uint32_t kw_bits(int kw) { return 0x09249249u << kw; }  // bits kw, kw+3, ...

// Accumulators (bit d) fed by at least one live box.
template <typename Params>
__device__ __forceinline__ uint32_t touched_accumulators(const Params& P, uint32_t live) {
  uint32_t t = 0;
  for (int p = 0; p < P.Din; ++p) {
    const PlaneInfo& pi = P.plane[p];
    if (pi.nkd != 0 && ((live >> (p * 3)) & 7)) t |= ((1u << pi.nkd) - 1u) << pi.dmin;
  }
  return t;
}

__device__ __forceinline__ uint32_t warp_or(uint32_t v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v |= __shfl_xor_sync(0xffffffffu, v, d);
  return v;
}

// ---------------------------------------------------------------------------------------------
// Plan: one block per (frame, tile row).  Phase 1 turns the flag bytes into per-tile live masks,
// phase 2 (one thread, <= 128 tiles) packs consecutive tiles into groups of <= 8 TMEM slots.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(64) k_conv3d_plan(const __grid_constant__ PlanParams Q) {
  __shared__ uint32_t s_live[kMaxTilesW];
  __shared__ uint8_t s_touched[kMaxTilesW], s_slot0[kMaxTilesW], s_restore[kMaxTilesW];
  __shared__ int32_t s_group[kMaxTilesW];  // group of the tile (-1: in no group) ...
  __shared__ uint8_t s_gpos[kMaxTilesW];   // ... and its position inside it
  const int b = blockIdx.x / Q.tiles_h, th = blockIdx.x % Q.tiles_h;
  const int tile_row0 = (b * Q.tiles_h + th) * Q.tiles_w;
  uint32_t valid = 0;
  for (int p = 0; p < Q.Din; ++p)
    if (Q.plane[p].nkd != 0) valid |= 7u << (p * 3);
  const uint32_t all_out = (1u << Q.Dout) - 1u;
  // rows 16*th-1 .. 16*th+16 of this tile row's boxes
  const bool ring_h = Q.force_ring && (th == 0 || (16 * th - 1 <= Q.H - 1 && Q.H - 1 <= 16 * th + 16));
  for (int tw = threadIdx.x; tw < Q.tiles_w; tw += blockDim.x) {
    uint32_t live = 0;
    if (Q.box_flags != nullptr) {
      for (int p = 0; p < Q.Din; ++p) {
        uint8_t* f = Q.box_flags + ((((size_t)b * Q.Din + p) * Q.tiles_h + th) * Q.tiles_w + tw) * 3;
        const uint32_t bits = (f[0] ? 1u : 0u) | (f[1] ? 2u : 0u) | (f[2] ? 4u : 0u);
        live |= bits << (p * 3);
        if (Q.consume && bits) { f[0] = 0; f[1] = 0; f[2] = 0; }
      }
      if (ring_h) live = ~0u;
      if (Q.force_ring && tw == 0) live |= kw_bits(0) | kw_bits(1);                // columns -1..6, 0..7 hold column 0
      if (Q.force_ring && tw == Q.tiles_w - 1) live |= kw_bits(1) | kw_bits(2);    // ... column W-1
    } else {
      live = ~0u;
    }
    live &= valid;
    if (Q.only_kw1) live &= kw_bits(1);
    const uint32_t touched = touched_accumulators(Q, live);
    uint32_t stale = all_out;
    if (Q.out_state != nullptr) {
      stale = Q.out_state[tile_row0 + tw];
      Q.out_state[tile_row0 + tw] = (uint8_t)touched;
    }
    s_live[tw] = live;
    s_group[tw] = -1;
    s_slot0[tw] = 0;
    s_touched[tw] = (uint8_t)touched;
    s_restore[tw] = (uint8_t)(stale & ~touched & all_out);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    // a group starts at a tile that has something to do, runs over at most kMaxGroupTiles tiles and
    // kSlots accumulators, and ends at its last tile that has something to do
    int start = -1, last = -1, cur_slots = 0;
    for (int tw = 0; tw < Q.tiles_w; ++tw) {
      const int n = __popc((uint32_t)s_touched[tw]);
      const bool active = n > 0 || s_restore[tw] != 0;
      if (start >= 0 && (cur_slots + n > kSlots || tw - start == kMaxGroupTiles)) {
        const int g = atomicAdd(&Q.hdr->n_groups, 1);
        Q.groups[g] = GroupDesc{tile_row0 + start, last - start + 1, b, (int16_t)th, (int16_t)start};
        for (int i = start; i <= last; ++i) { s_group[i] = g; s_gpos[i] = (uint8_t)(i - start); }
        start = -1;
      }
      if (start < 0) {
        if (!active) continue;
        start = tw;
        cur_slots = 0;
      }
      s_slot0[tw] = (uint8_t)cur_slots;
      cur_slots += n;
      if (active) last = tw;
    }
    if (start >= 0) {
      const int g = atomicAdd(&Q.hdr->n_groups, 1);
      Q.groups[g] = GroupDesc{tile_row0 + start, last - start + 1, b, (int16_t)th, (int16_t)start};
      for (int i = start; i <= last; ++i) { s_group[i] = g; s_gpos[i] = (uint8_t)(i - start); }
    }
  }
  __syncthreads();
  for (int tw = threadIdx.x; tw < Q.tiles_w; tw += blockDim.x) {
    TilePlan tp;
    tp.live = s_live[tw];
    tp.touched = s_touched[tw];
    tp.slot0 = s_slot0[tw];
    tp.nslots = (uint8_t)__popc((uint32_t)s_touched[tw]);
    tp.restore = s_restore[tw];
    Q.tiles[(size_t)tile_row0 + tw] = tp;
    if (s_group[tw] >= 0) Q.gtiles[(size_t)s_group[tw] * kMaxGroupTiles + s_gpos[tw]] = tp;
  }
}

// ---------------------------------------------------------------------------------------------
// The convolution.
// ---------------------------------------------------------------------------------------------
// kTrain = false: y = relu(acc*scale + shift) - sub (inference, BatchNorm folded).
// kTrain = true:  y = acc*scale + shift (the raw convolution: scale = 1, shift = bias) stored as bf16, and the
//                 per-channel sum / sum of squares of the fp32 values accumulated into P.stats - the batch
//                 statistics BatchNorm3d needs in training mode (model.py:153,161-162).
// kDebug = false is what every product call launches: no cycle counters, no experiment switches.
#define VN_TIMED(acc, stmt)                    \
  do {                                         \
    if constexpr (kDebug) {                    \
      const long long t_ = clock64();          \
      stmt;                                    \
      acc += clock64() - t_;                   \
    } else {                                   \
      stmt;                                    \
    }                                          \
  } while (0)

// kGather = true (the first layer of the eval fast path): there is no dense input tensor and no TMA box.  Warp 0 sends
//                 every box as a request to the filling warp of its A stage (warps 10..13), which re-zeroes what it
//                 wrote into the stage the last time, looks the 18 x 8 cells up in the voxelizer's occupancy bitmap,
//                 copies the few occupied rows (64 channels = 128 B each) from the compact (K,128) row table into the
//                 SWIZZLE_128B image the MMA expects, and arms the stage's full barrier.  tmap_x is unused.
// kS2 = true (the stride-2 heads of the RPN blocks, 2-D only): the tensor map samples every second column and row of the
//                 input; per (tile, chunk, kw) TWO boxes are loaded - 17 odd input rows 2 h0 - 1, + 1, ... serve kh = 0 (box rows
//                 0..15) and kh = 2 (box rows 1..16), 16 even rows 2 h0, ... serve kh = 1 - and each is its own message.
template <bool kTrain, bool kDebug, bool kGather, bool kS2 = false>
__global__ void __launch_bounds__(kGather ? kThreadsGather : kThreads, 1)
    k_conv3d_bn_relu(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ ConvParams P) {
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B operands need 1024-byte aligned tiles
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kNA * kABytes;
  Control* ctl = reinterpret_cast<Control*>(smem + kSmemData);

  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kNA; ++i) { mbar_init(&ctl->full_a[i], 1); mbar_init(&ctl->empty_a[i], 1); }
    for (int i = 0; i < kNB; ++i) { mbar_init(&ctl->full_b[i], 1); mbar_init(&ctl->empty_b[i], 1); }
    for (int i = 0; i < kSlots; ++i) { mbar_init(&ctl->acc_full[i], 1); mbar_init(&ctl->acc_empty[i], 8); }
    for (int i = 0; i < kSched; ++i) { mbar_init(&ctl->sched_full[i], 1); mbar_init(&ctl->sched_empty[i], 8); }
    if (kGather)
      for (int i = 0; i < kFill; ++i) { mbar_init(&ctl_mail(ctl)->full[i], 1); mbar_init(&ctl_mail(ctl)->empty[i], 1); }
    mbar_fence_init();
    fence_proxy_async_smem();
    if (!kGather) tma_prefetch_desc(&tmap_x);
  }
  if (kGather) {  // the A stages start all-zero and every filling warp restores that after each box
    if (threadIdx.x < kNA * 32) ctl_mail(ctl)->dirty[threadIdx.x >> 5][threadIdx.x & 31] = 0;
    if (threadIdx.x < kNA) ctl_mail(ctl)->turn[threadIdx.x] = 0;
    for (uint32_t i = threadIdx.x; i < kNA * kABytes / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem_a)[i] = make_uint4(0u, 0u, 0u, 0u);
    fence_proxy_async_smem();
  }
  if (threadIdx.x >= 64) {
    const int t = threadIdx.x - 64;  // 0..255
    if (t < 64) ctl->scale[t] = P.scale[t];
    if (kTrain && t < 128) ctl_stat(ctl)[t] = 0.0;
    const int nrow = (P.table_smem ? P.nvar : 1) * P.Dout;
    for (int i = t; i < nrow * 64; i += 256) ctl->table[i] = P.table_plain ? P.table[i & 63] : P.table[i];
    for (int i = t; i < P.Dout * 64; i += 256) ctl->sub[i] = P.sub != nullptr ? P.sub[i] : 0.0f;
    if (t < kSlots) {  // does the interior shift row of depth t cancel against `sub` exactly?
      bool z = t < P.Dout;
      if (z)
        for (int c = 0; c < 64; ++c) {
          const float sh = P.table_plain ? P.table[c] : P.table[t * 64 + c];
          const float sb = P.sub != nullptr ? P.sub[t * 64 + c] : 0.0f;
          z = z && (fmaxf(sh, 0.0f) - sb == 0.0f);
        }
      ctl->row_zero[t] = z ? 1u : 0u;
    }
  }
  if (warp == 1) {
    tmem_alloc(&ctl->tmem_base, kSlots * 64);
    tmem_relinquish();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = ctl->tmem_base;
  const int n_groups = P.hdr->n_groups;

  if (warp == 0) {
    // ====== scheduler + planner + TMA producer (warp-uniform; one elected lane issues) ======
    uint32_t a_it = 0, b_it = 0, s_it = 0;
    long long pw_a = 0, pw_b = 0, p_boxes = 0;
    const long long p_t0 = kDebug ? clock64() : 0;
    const int dbg = kDebug ? P.debug : 0;
    // The next group is fetched while the current one is processed: after the boxes of a group's first weight pass
    // are on their way, the work counter is bumped and the next group's descriptor and tile plans are loaded (ONE
    // dependent access: gtiles is indexed by the group); they are consumed when the group ends.  The ~1,500 clk of
    // memory latency sit where the 4-deep box ring hides them, not between two groups (measured before: ~4,000 clk
    // of idle tensor pipe per group switch - 3 dependent L2 accesses, then a cold first box).
    int g_raw = 0;
    if (lane == 0) g_raw = atomicAdd(&P.hdr->next_group, 1);
    int g = __shfl_sync(0xffffffffu, g_raw, 0);
    if (g >= n_groups) g = -1;
    else if (P.reverse) g = n_groups - 1 - g;  // groups in the opposite order of the layer that wrote the input: its last blocks are still in L2
    GroupDesc gd = GroupDesc{0, 0};
    TilePlan tp = TilePlan{0u, 0, 0, 0, 0};
    if (g >= 0) {
      gd = P.groups[g];
      if (lane < kMaxGroupTiles) tp = P.gtiles[(size_t)g * kMaxGroupTiles + lane];
      if ((int)lane >= gd.ntiles) tp = TilePlan{0u, 0, 0, 0, 0};
    }
    for (;;) {
      const uint32_t ss = s_it % kSched;
      mbar_wait(&ctl->sched_empty[ss], ((s_it / kSched) & 1) ^ 1);
      if (g >= 0 && lane < kMaxGroupTiles) ctl->sched_tiles[ss][lane] = tp;
      __syncwarp();
      if (lane == 0) {
        ctl->sched[ss] = g;
        ctl->sched_gd[ss] = gd;
        mbar_arrive(&ctl->sched_full[ss]);
      }
      ++s_it;
      if (g < 0) break;
      bool have_next = false;
      int g_next = -1;
      GroupDesc gd_next = GroupDesc{0, 0};
      TilePlan tp_next = TilePlan{0u, 0, 0, 0, 0};
      auto load_next = [&]() {
        if (lane == 0) g_raw = atomicAdd(&P.hdr->next_group, 1);
        g_next = __shfl_sync(0xffffffffu, g_raw, 0);
        if (g_next >= n_groups) g_next = -1;
        else if (P.reverse) g_next = n_groups - 1 - g_next;
        if (g_next >= 0) {
          gd_next = P.groups[g_next];
          if (lane < kMaxGroupTiles) tp_next = P.gtiles[(size_t)g_next * kMaxGroupTiles + lane];
        }
        have_next = true;
      };
      const uint32_t mylive = tp.live, mytouched = tp.touched, myslot0 = tp.slot0;
      uint32_t mystarted = 0;  // lane t: accumulators of tile t that already hold a partial sum
      const uint32_t any = warp_or(mylive);
      const int last_kw = (any & kw_bits(2)) ? 2 : ((any & kw_bits(1)) ? 1 : 0);
      const int tw0 = gd.tw0, th = gd.th, b = gd.b;
      const int h0 = th * kTileH;
      for (int c = 0; c < P.nchunk; ++c) {
        for (int kw = 0; kw < 3; ++kw) {
          if (!(any & kw_bits(kw))) continue;  // no tile of the group has a live box at this w shift
          const bool last_pass = (c == P.nchunk - 1) & (kw == last_kw);
          // all weights of this (chunk, kw): resident while the group's live boxes stream through
          const uint32_t sb = b_it % kNB;
          VN_TIMED(pw_b, mbar_wait(&ctl->empty_b[sb], ((b_it / kNB) & 1) ^ 1));
          if (elect_one()) {
            if (kDebug && (dbg & kDbgNoBLoad)) {
              mbar_arrive(&ctl->full_b[sb]);
            } else {
              mbar_arrive_expect_tx(&ctl->full_b[sb], kBSlotBytes);
              bulk_load_1d(smem_b + sb * kBSlotBytes, P.wpack + (size_t)(c * 3 + kw) * kBSlotBytes, kBSlotBytes,
                           &ctl->full_b[sb]);
            }
          }
          __syncwarp();
          ++b_it;
          uint32_t new_pass = kMsgNewPass;
          for (int t = 0; t < gd.ntiles; ++t) {
            uint32_t m = __shfl_sync(0xffffffffu, mylive, t) & kw_bits(kw);
            const uint32_t touched = __shfl_sync(0xffffffffu, mytouched, t);
            const uint32_t slot0 = __shfl_sync(0xffffffffu, myslot0, t);
            uint32_t started = __shfl_sync(0xffffffffu, mystarted, t);
            const int w0 = (tw0 + t) * kTileW;
            const bool tile_has_box = m != 0;
            while (m) {
              const int p = (__ffs(m) - 1) / 3;
              m &= m - 1;
              const PlaneInfo& pi = P.plane[p];
              const uint32_t range = ((1u << pi.nkd) - 1u) << pi.dmin;
              const uint32_t fresh = (range & ~started) >> pi.dmin;  // bit j: tap j starts its accumulator
              const uint32_t slot_lo = slot0 + __popc(touched & ((1u << pi.dmin) - 1u));
              const uint32_t full = (1u << pi.nkd) - 1u;
              // the tile's last box of the group's last weight pass also carries the "accumulators complete" bits (no
              // separate control message, which would take a ring slot away from the boxes in flight)
              const uint32_t done = (last_pass && m == 0) ? ((1u << __popc(touched)) - 1u) << slot0 : 0u;
              const uint32_t flags = (uint32_t)pi.nkd | new_pass | (fresh << kMsgFreshShift) | (done << kMsgCommitShift) |
                                     ((fresh << slot_lo) << kMsgWaitShift) | (fresh == 0 ? kMsgAccumulate : 0u) |
                                     ((fresh != 0 && fresh != full) ? kMsgSplit : 0u);
              const uint32_t sa = a_it % kNA;
              const uint32_t sf = a_it % kFill;  // kGather: the message's mailbox
              VN_TIMED(pw_a, mbar_wait(kGather ? &ctl_mail(ctl)->empty[sf] : &ctl->empty_a[sa], ((a_it / (kGather ? kFill : kNA)) & 1) ^ 1));
              if (kGather && lane < (uint32_t)kHaloH) {
                // the occupancy words the filling warp will look up, on their way into this SM's L1 while the request waits
                const int hh = h0 - 1 + (int)lane;
                if (hh >= 0 && hh < P.H) {
                  const int ww = w0 + kw - 1 < 0 ? 0 : w0 + kw - 1;
                  const size_t word = (size_t)b * P.g_words + ((uint32_t)((p * P.H + hh) * P.W + ww) >> 5);
                  prefetch_l1(P.g_bitmap + word);
                  prefetch_l1(P.g_prefix + word);
                }
              }
              if (kDebug) {
                ++p_boxes;
                if (P.trace != nullptr && blockIdx.x == 0 && lane == 0 && (int)a_it < P.trace_n) {
                  P.trace[(size_t)a_it * 4 + 0] = (unsigned long long)clock64();
                  P.trace[(size_t)a_it * 4 + 3] = flags;
                }
              }
              if (kS2 && P.pointwise) {
                // 1x1: the box starts at the tile's own first row, so box row 0 is kh = 1; the message has no other kh row
                if (elect_one()) {
                  ctl->msg[sa] = make_uint4(flags | (2u << kMsgKhSkipShift), tmem_base + slot_lo * 64, (pi.woff + pi.wstride) >> 4, 0u);
                  mbar_arrive_expect_tx(&ctl->full_a[sa], kABytes);
                  tma_load_5d(smem_a + sa * kABytes, &tmap_x, &ctl->full_a[sa], c * 64, w0 + kw - 1, h0, p, b);
                }
              } else if (kS2) {
                // odd rows: kh = 0 and (one box row on, weights two kh rows on) kh = 2; the tile-complete bits go with the second box
                const uint32_t commit_bits = flags & (0xffu << kMsgCommitShift);
                if (elect_one()) {
                  ctl->msg[sa] = make_uint4((flags & ~commit_bits) | (1u << kMsgKhSkipShift), tmem_base + slot_lo * 64, pi.woff >> 4,
                                            (2 * pi.wstride) >> 4);
                  mbar_arrive_expect_tx(&ctl->full_a[sa], kABytesS2);
                  tma_load_5d(smem_a + sa * kABytes, &tmap_x, &ctl->full_a[sa], c * 64, 2 * w0 + kw - 1, 2 * h0 - 1, p, b);
                }
                __syncwarp();
                ++a_it;
                const uint32_t sa2 = a_it % kNA;
                mbar_wait(&ctl->empty_a[sa2], ((a_it / kNA) & 1) ^ 1);
                if (elect_one()) {  // even rows: kh = 1
                  ctl->msg[sa2] = make_uint4((uint32_t)pi.nkd | kMsgAccumulate | commit_bits | (2u << kMsgKhSkipShift),
                                             tmem_base + slot_lo * 64, (pi.woff + pi.wstride) >> 4, 0u);
                  mbar_arrive_expect_tx(&ctl->full_a[sa2], kABytesS2);
                  tma_load_5d(smem_a + sa2 * kABytes, &tmap_x, &ctl->full_a[sa2], c * 64, 2 * w0 + kw - 1, 2 * h0, p, b);
                }
              } else if (elect_one()) {
                const uint4 msg = make_uint4(flags, tmem_base + slot_lo * 64, pi.woff >> 4, pi.wstride >> 4);
                if (kGather) {
                  ctl_mail(ctl)->req[sf] = make_int4(w0 + kw - 1, h0 - 1, p | (c << 8) | (kReqBox << 16), b);
                  ctl_mail(ctl)->msg[sf] = msg;
                  mbar_arrive(&ctl_mail(ctl)->full[sf]);
                } else if (kDebug && (dbg & kDbgNoALoad)) {
                  ctl->msg[sa] = msg;
                  mbar_arrive(&ctl->full_a[sa]);
                } else {
                  ctl->msg[sa] = msg;
                  mbar_arrive_expect_tx(&ctl->full_a[sa], kABytes);
                  tma_load_5d(smem_a + sa * kABytes, &tmap_x, &ctl->full_a[sa], c * 64, w0 + kw - 1, h0 - 1, p, b);
                }
              }
              __syncwarp();
              started |= range;
              new_pass = 0;
              ++a_it;
            }
            if ((int)lane == t) mystarted = started;
            // last weight pass of the group: tile t has received everything it will get (no box of its own in this pass)
            if (last_pass && touched != 0 && !tile_has_box) {
              const uint32_t done = ((1u << __popc(touched)) - 1u) << slot0;
              const uint32_t sa = a_it % kNA;
              const uint32_t sf = a_it % kFill;
              mbar_wait(kGather ? &ctl_mail(ctl)->empty[sf] : &ctl->empty_a[sa], ((a_it / (kGather ? kFill : kNA)) & 1) ^ 1);
              if (elect_one()) {
                const uint4 msg = make_uint4(done << kMsgCommitShift, 0u, 0u, 0u);
                if (kGather) {
                  ctl_mail(ctl)->req[sf] = make_int4(0, 0, kReqControl << 16, 0);
                  ctl_mail(ctl)->msg[sf] = msg;
                  mbar_arrive(&ctl_mail(ctl)->full[sf]);
                } else {
                  ctl->msg[sa] = msg;
                  mbar_arrive(&ctl->full_a[sa]);
                }
              }
              __syncwarp();
              ++a_it;
            }
          }
          if (!have_next) load_next();  // (after the first pass: the counter bump has returned by now)
        }
      }
      if (!have_next) load_next();
      if ((int)lane >= gd_next.ntiles) tp_next = TilePlan{0u, 0, 0, 0, 0};
      g = g_next;
      gd = gd_next;
      tp = tp_next;
    }
    {  // tell the issuing warp to stop
      const uint32_t sa = a_it % kNA;
      const uint32_t sf = a_it % kFill;
      mbar_wait(kGather ? &ctl_mail(ctl)->empty[sf] : &ctl->empty_a[sa], ((a_it / (kGather ? kFill : kNA)) & 1) ^ 1);
      if (elect_one()) {
        const uint4 msg = make_uint4(kMsgEnd, 0u, 0u, 0u);
        if (kGather) {
          ctl_mail(ctl)->req[sf] = make_int4(0, 0, kReqControl << 16, 0);
          ctl_mail(ctl)->msg[sf] = msg;
          mbar_arrive(&ctl_mail(ctl)->full[sf]);
        } else {
          ctl->msg[sa] = msg;
          mbar_arrive(&ctl->full_a[sa]);
        }
      }
      __syncwarp();
      ++a_it;
    }
    if (kGather) {  // ... and then every filling warp (each sees its requests in order)
      for (int i = 0; i < kFill; ++i, ++a_it) {
        const uint32_t sf = a_it % kFill;
        mbar_wait(&ctl_mail(ctl)->empty[sf], ((a_it / kFill) & 1) ^ 1);
        if (elect_one()) {
          ctl_mail(ctl)->req[sf] = make_int4(0, 0, kReqStop << 16, 0);
          mbar_arrive(&ctl_mail(ctl)->full[sf]);
        }
        __syncwarp();
      }
    }
    if (kDebug && P.prof != nullptr && lane == 0) {
      unsigned long long* o = P.prof + (size_t)blockIdx.x * 16;
      o[0] = (unsigned long long)(clock64() - p_t0); o[1] = pw_a; o[2] = pw_b; o[3] = p_boxes;
    }
  } else if (warp == 1) {
    // ============ MMA issuer: ONE thread interprets the producer's messages ============
    // Software-pipelined.  Measured with the instrumented instantiation (tools/prof_conv.py, tools/probes/mma_probe.cu):
    // the tensor pipe accepts only ~2 MMAs ahead of the one it executes (a 128x64x16 MMA takes 48 clk, 128x128x16 64 clk),
    // so whatever the issuing thread does between two MMAs that takes longer than that leaves the pipe idle - and it used
    // to spend ~300 clk between the last MMA of one message and the first of the next (warp-wide waits, elections, the
    // message read).  Now a single thread runs the loop (no warp collectives), and the next message is probed, read and
    // decoded BETWEEN the MMAs of the current one; only what may legitimately block (a new weight pass, an accumulator
    // the epilogue still holds) stays in front of a message's first MMA.
    long long mw_a = 0, mw_b = 0, mw_acc = 0;
    const long long m_t0 = kDebug ? clock64() : 0;
    if (elect_one()) {
      uint32_t a_it = 0, b_it = 0;
      uint32_t acc_phase = 0;  // bit s: parity of the number of times TMEM slot s has been (re)started
      const uint32_t a_lo0 = ((smem_u32(smem_a) >> 4) & 0x3FFFu) | (1u << 16);
      const uint32_t b_lo0 = ((smem_u32(smem_b) >> 4) & 0x3FFFu) | (1u << 16);
      uint32_t b_lo_pass = b_lo0;
      const bool no_mma = kDebug && (P.debug & kDbgNoMma);
      // decoded message: everything the MMAs of a box need, computed as soon as the message has been read
      struct Dec {
        uint32_t f, d_tmem, a_lo, bz, bw, idesc, nkd;
      };
      auto decode = [&](const uint4& m, uint32_t it) {
        Dec d;
        d.f = m.x;
        d.nkd = m.x & kMsgNkdMask;
        d.d_tmem = m.y;
        d.a_lo = a_lo0 + (it % kNA) * (kABytes >> 4);
        d.bz = m.z;
        d.bw = m.w;
        d.idesc = umma_idesc_bf16(128, 64) + ((d.nkd - 1u) << 20);  // N = 64 * nkd
        return d;
      };
      VN_TIMED(mw_a, mbar_wait(&ctl->full_a[0], 0));
      Dec cur = decode(lds_u4(&ctl->msg[0]), 0);
      for (;;) {
        const uint32_t sa = a_it % kNA;
        const bool tr = kDebug && P.trace != nullptr && blockIdx.x == 0 && (int)a_it < P.trace_n;
        if (tr) P.trace[(size_t)a_it * 4 + 1] = (unsigned long long)clock64();
        const uint32_t f = cur.f;
        if (f & (kMsgNewPass | (0xffu << kMsgWaitShift))) {  // (rare) things that may block in front of the MMAs
          if (f & kMsgNewPass) {
            // the previous pass is fully issued: its weights may be overwritten once it completes
            if (b_it != 0) umma_commit(&ctl->empty_b[(b_it - 1) % kNB]);
            VN_TIMED(mw_b, mbar_wait(&ctl->full_b[b_it % kNB], (b_it / kNB) & 1));
            b_lo_pass = b_lo0 + (b_it % kNB) * (kBSlotBytes >> 4);
            ++b_it;
          }
          const uint32_t wait_mask = (f >> kMsgWaitShift) & 0xffu;
          uint32_t wm = wait_mask;
          while (wm) {  // accumulators started by this box must have been drained by the epilogue
            const uint32_t s = __ffs(wm) - 1;
            wm &= wm - 1;
            VN_TIMED(mw_acc, mbar_wait(&ctl->acc_empty[s], ((acc_phase >> s) & 1) ^ 1));
          }
          acc_phase ^= wait_mask;
        }
        tc_fence_after_sync();
        const uint32_t sn = (a_it + 1) % kNA, pn = ((a_it + 1) / kNA) & 1;
        bool have_next = false;
        Dec nxt = cur;
        if (cur.nkd != 0) {
          const uint32_t d_tmem = cur.d_tmem, a_lo = cur.a_lo, idesc = cur.idesc;
          const uint32_t b_lo = b_lo_pass + cur.bz;
          if (!no_mma) {  // kh = 0
            if (f & kMsgSplit) {  // first k-step per accumulator (N = 64): accumulate only where a partial sum exists
              const uint32_t fresh = (f >> kMsgFreshShift) & 7u;
              for (uint32_t j = 0; j < cur.nkd; ++j)
                umma_bf16_ss(d_tmem + j * 64, desc_from_lo(a_lo), desc_from_lo(b_lo + j * (kBTapBytes >> 4)),
                             umma_idesc_bf16(128, 64), ((fresh >> j) & 1) ^ 1);
            } else {
              umma_bf16_ss(d_tmem, desc_from_lo(a_lo), desc_from_lo(b_lo), idesc, (f & kMsgAccumulate) ? 1u : 0u);
            }
#pragma unroll
            for (int k = 1; k < 4; ++k) umma_bf16_ss(d_tmem, desc_from_lo(a_lo + k * 2), desc_from_lo(b_lo + k * 2), idesc, 1u);
          }
          // in the shadow of those MMAs: is the next message there already?  (normally yes: the producer runs ahead)
          // Read and decode it now, so that after this box's last MMA only the uniform-register moves are left.
          have_next = mbar_test_wait(&ctl->full_a[sn], pn);
          if (have_next) nxt = decode(lds_u4(&ctl->msg[sn]), a_it + 1);
          if (!no_mma) {  // kh = 1, 2
            const int n_extra = kS2 ? 2 - (int)((f >> kMsgKhSkipShift) & 3u) : 2;
#pragma unroll
            for (int kh = 1; kh < 3; ++kh) {
              if (kS2 && kh > n_extra) break;
              const uint32_t a = a_lo + kh * ((kTileW * 128) >> 4);
              const uint32_t bq = b_lo + kh * cur.bw;
#pragma unroll
              for (int k = 0; k < 4; ++k) umma_bf16_ss(d_tmem, desc_from_lo(a + k * 2), desc_from_lo(bq + k * 2), idesc, 1u);
            }
          }
          umma_commit(&ctl->empty_a[sa]);
          uint32_t cm = (f >> kMsgCommitShift) & 0xffu;  // (tile complete after this box)
          while (cm) {
            const uint32_t s = __ffs(cm) - 1;
            cm &= cm - 1;
            umma_commit(&ctl->acc_full[s]);
          }
        } else {
          uint32_t cm = (f >> kMsgCommitShift) & 0xffu;
          while (cm) {
            const uint32_t s = __ffs(cm) - 1;
            cm &= cm - 1;
            umma_commit(&ctl->acc_full[s]);
          }
          mbar_arrive(&ctl->empty_a[sa]);
        }
        if (tr) P.trace[(size_t)a_it * 4 + 2] = (unsigned long long)clock64();
        ++a_it;
        if (f & kMsgEnd) break;
        if (!have_next) {
          VN_TIMED(mw_a, mbar_wait(&ctl->full_a[sn], pn));
          nxt = decode(lds_u4(&ctl->msg[sn]), a_it);
        }
        cur = nxt;
      }
      if (kDebug && P.prof != nullptr) {
        unsigned long long* o = P.prof + (size_t)blockIdx.x * 16;
        o[4] = (unsigned long long)(clock64() - m_t0); o[5] = mw_a; o[6] = mw_b; o[7] = mw_acc;
      }
    }
    __syncwarp();
  } else if (kGather && warp >= 10) {
    // ===================== box filling (warps 10 .. 10 + kFill - 1; warp 10 + f takes messages f, f + kFill, ...) =====================
    // Box row c = 8 * (image row) + column, 0..143, belongs to lane c % 32 (five candidates per lane).  Per request:
    // look the 18 x 8 cells up (occupancy words + popcount prefix: ONE round of independent L2 accesses), load the
    // occupied rows of this lane into registers (a second round), and only then wait for the stage - both latencies
    // sit behind the MMAs that still read the stage.  Then: re-zero what this lane wrote the last time, store the new
    // rows in the SWIZZLE_128B image (16-byte chunk q of row c at q ^ (c & 7)), publish the issuer message, arm full_a.
    const int f = (int)warp - 10, j = f % kNA;  // mailbox, stage (with kFill = 2 kNA two warps use a stage alternately)
    Mailboxes* mail = ctl_mail(ctl);
    uint8_t* stage = smem_a + j * kABytes;
    long long total = P.g_voxoff[P.B];
    if (total > P.g_max_voxels) total = P.g_max_voxels;
    uint32_t n_use = 0;  // messages this warp has taken
    long long fw_req = 0, f_look = 0, fw_empty = 0, f_store = 0, f_boxes = 0;
    for (;;) {
      VN_TIMED(fw_req, mbar_wait(&mail->full[f], n_use & 1));
      const long long f_t1 = kDebug ? clock64() : 0;
      const uint4 rq_u = lds_u4(&mail->req[f]);
      const uint4 msg = lds_u4(&mail->msg[f]);
      __syncwarp();
      if (lane == 0) mbar_arrive(&mail->empty[f]);
      const int4 rq = make_int4((int)rq_u.x, (int)rq_u.y, (int)rq_u.z, (int)rq_u.w);
      const int kind = rq.z >> 16;
      if (kind == kReqStop) break;
      uint32_t mine = 0;
      int row_of[5] = {0, 0, 0, 0, 0};
      uint4 v0[8], v1[8];
      int k0 = -1, k1 = -1;
      const __nv_bfloat16* src0 = P.g_rows + ((rq.z >> 8) & 0xff) * 64;
      if (kind == kReqBox) {
        // lanes 0..17: occupancy bits of image row rq.y + lane, columns rq.x .. rq.x + 7, and the table row of the first
        uint32_t mask = 0, first = 0;
        const int h = rq.y + (int)lane;
        if (lane < (uint32_t)kHaloH && h >= 0 && h < P.H && !(kDebug && (P.debug & kDbgNoLookup))) {
          const int lo_w = rq.x < 0 ? 0 : rq.x;
          const int hi_w = rq.x + kTileW > P.W ? P.W : rq.x + kTileW;
          const int nb = hi_w - lo_w;
          const uint32_t cell = (uint32_t)(((rq.z & 0xff) * P.H + h) * P.W + lo_w);
          const uint32_t word = cell >> 5, sh = cell & 31u;
          const uint32_t* bm = P.g_bitmap + (size_t)rq.w * P.g_words + word;
          const uint32_t w_lo = __ldg(bm);
          const uint32_t w_hi = sh + (uint32_t)nb > 32u ? __ldg(bm + 1) : 0u;
          const uint32_t pre = __ldg(P.g_prefix + (size_t)rq.w * P.g_words + word);
          const uint32_t off = (uint32_t)__ldg(P.g_voxoff + rq.w);
          const uint32_t win = __funnelshift_r(w_lo, w_hi, sh) & ((1u << nb) - 1u);
          mask = win << (lo_w - rq.x);
          first = off + pre + __popc(w_lo & ((1u << sh) - 1u));
        }
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          const uint32_t c = lane + 32 * k;
          const uint32_t r = c < 144u ? c >> 3 : 17u, i = c & 7u;
          const uint32_t m_r = __shfl_sync(0xffffffffu, mask, r);
          const uint32_t f_r = __shfl_sync(0xffffffffu, first, r);
          const long long row = (long long)f_r + __popc(m_r & ((1u << i) - 1u));
          if (c < 144u && ((m_r >> i) & 1u) && row < total) {
            mine |= 1u << k;
            row_of[k] = (int)row;
          }
        }
        if (kDebug) f_look += clock64() - f_t1 + (mine & 0u);
        // the first two occupied candidates of this lane travel while the stage is still being read
        uint32_t rest = mine;
        if (rest) { k0 = __ffs(rest) - 1; rest &= rest - 1; }
        if (rest) { k1 = __ffs(rest) - 1; }
        if (k0 >= 0) {
          const int r0 = k0 == 0 ? row_of[0] : (k0 == 1 ? row_of[1] : (k0 == 2 ? row_of[2] : (k0 == 3 ? row_of[3] : row_of[4])));
          const uint4* src = reinterpret_cast<const uint4*>(src0 + (size_t)r0 * 128);
#pragma unroll
          for (int q = 0; q < 8; ++q) v0[q] = __ldg(src + q);
        }
        if (k1 >= 0) {
          const int r1 = k1 == 1 ? row_of[1] : (k1 == 2 ? row_of[2] : (k1 == 3 ? row_of[3] : row_of[4]));
          const uint4* src = reinterpret_cast<const uint4*>(src0 + (size_t)r1 * 128);
#pragma unroll
          for (int q = 0; q < 8; ++q) v1[q] = __ldg(src + q);
        }
      }
      // This is use number u = (kFill / kNA) n + f / kNA of the stage.  A parity wait can only tell "the previous phase is over" from
      // "it is not", and this warp could otherwise get here before the other warp of the stage has even armed use u - 1
      // (two phases back reads as over): first wait until use u - 1 is armed, then until its MMAs are complete.
      const uint32_t use = (uint32_t)(kFill / kNA) * n_use + (uint32_t)(f / kNA);
      {
        const long long t_ = kDebug ? clock64() : 0;
        while (*reinterpret_cast<volatile uint32_t*>(&mail->turn[j]) != use) {
        }
        mbar_wait(&ctl->empty_a[j], (use & 1) ^ 1);
        if (kDebug) fw_empty += clock64() - t_;
      }
      ++n_use;
      const long long f_t4 = kDebug ? clock64() : 0;
      if (kind == kReqBox && !(kDebug && (P.debug & kDbgNoALoad))) {
        const uint32_t prev = mail->dirty[j][lane];
#pragma unroll
        for (int k = 0; k < 5; ++k)
          if ((prev >> k) & 1) {
            uint4* dst = reinterpret_cast<uint4*>(stage + (lane + 32 * k) * 128);
#pragma unroll
            for (int q = 0; q < 8; ++q) dst[q] = make_uint4(0u, 0u, 0u, 0u);
          }
        if (k0 >= 0) {
          const uint32_t c = lane + 32 * k0;
#pragma unroll
          for (int q = 0; q < 8; ++q) *reinterpret_cast<uint4*>(stage + c * 128 + ((q ^ (c & 7u)) << 4)) = v0[q];
        }
        if (k1 >= 0) {
          const uint32_t c = lane + 32 * k1;
#pragma unroll
          for (int q = 0; q < 8; ++q) *reinterpret_cast<uint4*>(stage + c * 128 + ((q ^ (c & 7u)) << 4)) = v1[q];
        }
        const uint32_t more = mine & ~((k0 >= 0 ? 1u << k0 : 0u) | (k1 >= 0 ? 1u << k1 : 0u));
        if (more) {  // (rare) a third .. fifth occupied candidate
#pragma unroll
          for (int k = 2; k < 5; ++k)
            if ((more >> k) & 1) {
              const uint4* src = reinterpret_cast<const uint4*>(src0 + (size_t)row_of[k] * 128);
              const uint32_t c = lane + 32 * k;
#pragma unroll
              for (int q = 0; q < 8; ++q) v0[q] = __ldg(src + q);
#pragma unroll
              for (int q = 0; q < 8; ++q) *reinterpret_cast<uint4*>(stage + c * 128 + ((q ^ (c & 7u)) << 4)) = v0[q];
            }
        }
        mail->dirty[j][lane] = (uint8_t)mine;
        if (kDebug) { f_store += clock64() - f_t4; f_boxes -= clock64(); }
        if (!(kDebug && (P.debug & kDbgNoFence))) fence_proxy_async_smem();  // generic-proxy stores -> visible to the tensor core's operand reads
        if (kDebug) f_boxes += clock64();
      }
      if (lane == 0) ctl->msg[j] = msg;
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&ctl->full_a[j]);
        *reinterpret_cast<volatile uint32_t*>(&mail->turn[j]) = use + 1u;
      }
    }
    if (kDebug && P.prof != nullptr && f == 0 && lane == 0) {
      unsigned long long* o = P.prof + (size_t)blockIdx.x * 16;
      o[11] = fw_req; o[12] = f_look; o[13] = fw_empty; o[14] = f_store; o[15] = f_boxes;
    }
  } else {
    // ===================== epilogue (warps 2..9) =====================
    // Two warps per TMEM lane quadrant (a warp may only read lanes 32*(warp%4)..+31): warps 2..5 take
    // output channels 0..31 of their 32 positions, warps 6..9 channels 32..63.
    const uint32_t quad = warp & 3;
    const uint32_t half = (warp - 2) >> 2;
    const uint32_t m = quad * 32 + lane;           // accumulator row = output position in the tile
    const int hl = m / kTileW, wl = m % kTileW;
    const float* scale_h = ctl->scale + half * 32;
    const float relu_floor = P.relu_floor;  // 0 (ReLU), or -inf for a layer without activation (the 1x1 heads of the RPN)
    uint32_t s_it = 0, acc_phase = 0;
    float st_s[kTrain ? 32 : 1], st_q[kTrain ? 32 : 1];
#pragma unroll
    for (int j = 0; j < (kTrain ? 32 : 1); ++j) { st_s[j] = 0.0f; st_q[j] = 0.0f; }
    long long ew_sched = 0, ew_acc = 0;
    const long long e_t0 = kDebug ? clock64() : 0;
    const bool no_store = kDebug && (P.debug & kDbgNoStore);
    for (;;) {
      const uint32_t ss = s_it % kSched;
      VN_TIMED(ew_sched, mbar_wait(&ctl->sched_full[ss], (s_it / kSched) & 1));
      const int g = ctl->sched[ss];
      ++s_it;
      if (g < 0) break;
      const GroupDesc gd = ctl->sched_gd[ss];  // (the ring slot is handed back when the group is done, below)
      const int tw0 = gd.tw0, th = gd.th, b = gd.b;
      const int h = th * kTileH + hl;
      const bool inside = h < P.H;
      // border class of the output row: 0 interior, 1 first, 2 last (and with 25 variants 3 second, 4 last but one)
      const int vh = h == 0 ? 1 : (h == P.H - 1 ? 2 : (P.nvar == 25 ? (h == 1 ? 3 : (h == P.H - 2 ? 4 : 0)) : 0));
      // tile rows 1..H-2 (2..H-3) only, all of them inside the image: every position is of the interior class
      const bool rows_interior = th > 0 && (th + 1) * kTileH + (P.nvar == 25 ? 1 : 0) < P.H;
      uint32_t used = 0;
      for (int t = 0; t < gd.ntiles; ++t) {
        const TilePlan tp = ctl->sched_tiles[ss][t];
        const uint32_t touched = tp.touched;
        used |= ((1u << tp.nslots) - 1u) << tp.slot0;
        const int tw = tw0 + t;
        const int w = tw * kTileW + wl;
        const int vw = w == 0 ? 1 : (w == P.W - 1 ? 2 : (P.nvar == 25 ? (w == 1 ? 3 : (w == P.W - 2 ? 4 : 0)) : 0));
        const int variant = P.nvar == 9 ? vh * 3 + vw : (P.nvar == 25 ? vh * 5 + vw : 0);
        const bool tile_interior = P.nvar == 1 ? (th + 1) * kTileH <= P.H : (rows_interior && tw > 0 && tw < P.tiles_w - 1);
        for (int dout = 0; dout < P.Dout; ++dout) {
          const bool has_acc = (touched >> dout) & 1;  // else: no live box reached this accumulator
          if (!has_acc && !((tp.restore >> dout) & 1)) continue;  // the block already holds the layer background
          if (!kTrain && !has_acc && P.out_mode == 0 && tile_interior && ctl->row_zero[dout]) {
            if (no_store) continue;
            // nothing but background in reach and the layer stores deviations: the 16 KB block is zero.
            // This warp clears tile rows 4*quad + 2*half, +1 (1 KB each, contiguous).
            __nv_bfloat16* dst0 = reinterpret_cast<__nv_bfloat16*>(P.out) +
                                  ((((size_t)b * P.Dout + dout) * P.H + th * kTileH + quad * 4 + half * 2) * P.W + tw * kTileW) * 64;
#pragma unroll
            for (int r = 0; r < 2; ++r) {
              uint4* dst = reinterpret_cast<uint4*>(dst0 + (size_t)r * P.W * 64);
              dst[lane] = make_uint4(0u, 0u, 0u, 0u);
              dst[lane + 32] = make_uint4(0u, 0u, 0u, 0u);
            }
            continue;
          }
          const uint32_t slot = tp.slot0 + __popc(touched & ((1u << dout) - 1u));
          uint32_t r[32];
          if (has_acc) {
            VN_TIMED(ew_acc, mbar_wait(&ctl->acc_full[slot], (acc_phase >> slot) & 1));
            tc_fence_after_sync();
            tmem_ld_32x32b_x32(tmem_base + ((quad * 32) << 16) + slot * 64 + half * 32, r);
            tmem_ld_wait();
            // accumulator read: hand it back to the MMA issuer before the global stores
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl->acc_empty[slot]);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) r[j] = 0u;
          }
          if (inside && w < P.W && !no_store) {  // (w >= W: the 2-D layers allow a ragged last tile column)
            const float* shift_row = (variant == 0 || P.table_smem) ? ctl->table + (variant * P.Dout + dout) * 64 + half * 32
                                                                    : P.table + (variant * P.Dout + dout) * 64 + half * 32;
            const float* sub_row = (variant == 0 || P.sub_classes == 1) ? ctl->sub + dout * 64 + half * 32
                                                                         : P.sub + (variant * P.Dout + dout) * 64 + half * 32;
            float v[32];
            if (kTrain) {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                v[j] = fmaf(__uint_as_float(r[j]), scale_h[j], shift_row[j]);
                st_s[j] += v[j];
                st_q[j] = fmaf(v[j], v[j], st_q[j]);
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                v[j] = fmaxf(fmaf(__uint_as_float(r[j]), scale_h[j], shift_row[j]), relu_floor) - sub_row[j];
            }
            if (P.out_mode != 1) {
              __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(P.out) +
                                   (P.out_mode == 0 ? ((((size_t)b * P.Dout + dout) * P.H + h) * P.W + w) * 64 + half * 32
                                                    : (((size_t)b * (P.H * P.out_up) + h * P.out_up + P.out_oy) * (P.W * P.out_up) + w * P.out_up + P.out_ox) * P.out_ctotal +
                                                          P.out_coff + dout * 64 + half * 32);
#pragma unroll
              for (int q = 0; q < 2; ++q) {  // 2 x 32 B: one full sector per lane per store
                uint32_t pk[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                  const __nv_bfloat162 t2 = __floats2bfloat162_rn(v[q * 16 + 2 * e], v[q * 16 + 2 * e + 1]);
                  pk[e] = *reinterpret_cast<const uint32_t*>(&t2);
                }
                st_global_v8(dst + q * 16, pk);
              }
            } else {
              const size_t plane = (size_t)P.H * P.W;
              float* dst = reinterpret_cast<float*>(P.out) +
                           ((size_t)b * 64 * P.Dout + (size_t)(half * 32) * P.Dout + dout) * plane + (size_t)h * P.W + w;
#pragma unroll
              for (int j = 0; j < 32; ++j) st_global_cs_f32(dst + (size_t)j * P.Dout * plane, v[j]);  // written once, read by another kernel: do not displace the input boxes in L2
            }
          }
        }
      }
      acc_phase ^= used;
      __syncwarp();
      if (lane == 0) mbar_arrive(&ctl->sched_empty[ss]);
    }
    if (kDebug && P.prof != nullptr && threadIdx.x == 64) {
      unsigned long long* o = P.prof + (size_t)blockIdx.x * 16;
      o[8] = (unsigned long long)(clock64() - e_t0); o[9] = ew_sched; o[10] = ew_acc;
    }
    if (kTrain) {
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        atomicAdd(&ctl_stat(ctl)[half * 32 + j], (double)st_s[j]);
        atomicAdd(&ctl_stat(ctl)[64 + half * 32 + j], (double)st_q[j]);
      }
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after_sync();
    tmem_dealloc(tmem_base, kSlots * 64);
  }
  if (kTrain && threadIdx.x >= 64 && threadIdx.x < 64 + 128) atomicAdd(&P.stats[threadIdx.x - 64], ctl_stat(ctl)[threadIdx.x - 64]);
}

// Weight image.  Depth taps are grouped into "kd classes" that can be valid for the same input plane:
//   stride_d 1: one class {2,1,0};  stride_d 2: class 0 = {2,0} (planes with p+pad even), class 1 = {1}.
// Layout: [chunk][kw][class][kh][j: kd descending][n = 64][k = 64] bf16, rows 128B-swizzled: the 72 KB
// of one (chunk, kw) are contiguous (one bulk copy), and the taps valid for a plane sit side by side.
struct KdClasses {
  int n_classes;
  int nkd[2];
  int kd[2][3];
  uint32_t base[2];  // byte offset of the class inside a (chunk, kw) group
};

__host__ __device__ inline KdClasses kd_classes(int stride_d) {
  KdClasses k{};
  if (stride_d == 1) {
    k.n_classes = 1;
    k.nkd[0] = 3; k.kd[0][0] = 2; k.kd[0][1] = 1; k.kd[0][2] = 0;
    k.base[0] = 0;
  } else {
    k.n_classes = 2;
    k.nkd[0] = 2; k.kd[0][0] = 2; k.kd[0][1] = 0;
    k.nkd[1] = 1; k.kd[1][0] = 1;
    k.base[0] = 0;
    k.base[1] = 3 * 2 * kBTapBytes;  // after class 0: 3 kh x 2 taps
  }
  return k;
}

// torch [64][cin][3][3][3] fp32 -> the image above.
__global__ void __launch_bounds__(256)
    k_pack_weights(const float* __restrict__ w, int cin, int stride_d, uint8_t* __restrict__ packed) {
  const int nchunk = cin / 64;
  const KdClasses kc = kd_classes(stride_d);
  const int total = 27 * cin * 64;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    int r = i;
    const int k = r % 64; r /= 64;
    const int n = r % 64; r /= 64;
    const int kh = r % 3; r /= 3;
    const int kw = r % 3; r /= 3;
    const int c = r % nchunk; r /= nchunk;
    const int kd = r;
    int cls = 0, j = 0;
    for (int a = 0; a < kc.n_classes; ++a)
      for (int q = 0; q < kc.nkd[a]; ++q)
        if (kc.kd[a][q] == kd) { cls = a; j = q; }
    const float v = w[((((size_t)n * cin + (c * 64 + k)) * 3 + kd) * 3 + kh) * 3 + kw];
    const size_t tile = (size_t)(c * 3 + kw) * kBSlotBytes + kc.base[cls] + (size_t)(kh * kc.nkd[cls] + j) * kBTapBytes;
    const size_t off = tile + (size_t)n * 128 + ((((k >> 3) ^ (n & 7)) << 4) | ((k & 7) << 1));
    *reinterpret_cast<__nv_bfloat16*>(packed + off) = __float2bfloat16_rn(v);
  }
}

// Background folding.  When the layer's input is stored as a deviation from a per-(plane, channel)
// background bg_in (what the previous layer outputs wherever ITS input was all background), the
// background's own contribution to every output is a constant that depends only on the output depth
// and on which taps fall outside the image (the reference zero-pads, so out-of-image taps see 0, not
// the background): 3 h-classes (interior, h = 0, h = H-1) x 3 w-classes.  This kernel evaluates
//   bias[v][d'][n] = sum over in-image taps (kd, kh, kw), channels c of bf16(W[n][c][kd][kh][kw]) * bg_in[p][c]
// and writes the epilogue's shift table  table[v][d'][n] = bias * scale[n] + shift[n]  plus the layer's
// own output background per border class  bg_out[v][d'][n] = relu(table[v][d'][n]).
// One block per (variant, d'), one thread per output channel.
__global__ void __launch_bounds__(64)
    k_fold_background(const float* __restrict__ w, int cin, int Din, int Dout, int stride_d, int pad_d,
                      const float* __restrict__ scale, const float* __restrict__ shift,
                      const float* __restrict__ bg_in, int bg_classes, float* __restrict__ table, float* __restrict__ bg_out) {
  // bg_classes 1: bg_in [Din][cin], the same at every position -> 3 x 3 output variants (interior / first / last row, column).
  // bg_classes 9: bg_in [3][3][Din][cin] by the border class of the INPUT position (what a layer with 3 x 3 variants
  //               outputs over empty space) -> 5 x 5 output variants: interior, first, last, second, last but one row
  //               (column) - an output in row 1 still sees input row 0, whose background is the "first row" one.
  const int nv = bg_classes == 9 ? 5 : 3;
  const int v = blockIdx.x / Dout, d = blockIdx.x % Dout, n = threadIdx.x;
  const int vh = v / nv, vw = v % nv;
  double acc = 0.0;
  if (bg_in != nullptr) {
    for (int kd = 0; kd < 3; ++kd) {
      const int p = d * stride_d - pad_d + kd;
      if (p < 0 || p >= Din) continue;
      for (int kh = 0; kh < 3; ++kh) {
        if ((vh == 1 && kh == 0) || (vh == 2 && kh == 2)) continue;  // tap outside the image
        // border class of the input row this tap reads: output row 0 / 1 read input row 0 with kh = 1 / 0, ...
        const int ch = ((vh == 1 && kh == 1) || (vh == 3 && kh == 0)) ? 1 : (((vh == 2 && kh == 1) || (vh == 4 && kh == 2)) ? 2 : 0);
        for (int kw = 0; kw < 3; ++kw) {
          if ((vw == 1 && kw == 0) || (vw == 2 && kw == 2)) continue;
          const int cw = ((vw == 1 && kw == 1) || (vw == 3 && kw == 0)) ? 1 : (((vw == 2 && kw == 1) || (vw == 4 && kw == 2)) ? 2 : 0);
          const float* bg = bg_in + (bg_classes == 9 ? (size_t)(ch * 3 + cw) * Din * cin : 0);
          for (int c = 0; c < cin; ++c) {
            const float wv = __bfloat162float(__float2bfloat16_rn(w[((((size_t)n * cin + c) * 3 + kd) * 3 + kh) * 3 + kw]));
            acc += (double)wv * (double)bg[p * cin + c];
          }
        }
      }
    }
  }
  const float t = (float)(acc * (double)scale[n] + (double)shift[n]);
  table[((size_t)v * Dout + d) * 64 + n] = t;
  if (bg_out != nullptr) bg_out[((size_t)v * Dout + d) * 64 + n] = fmaxf(t, 0.0f);  // (3 x 3 variants only: checked by the host)
}

// Voxel coordinates -> live-box flags of a chain of up to three stacked 3x3x3 layers.  A voxel at
// (z, y, x) makes the input of layer l deviate from its background at most at planes S_l (S_0 = {z},
// S_{l+1} = output depths fed by S_l), rows y-l..y+l and columns x-l..x+l; every box (plane, th, tw, kw)
// of layer l that contains such a position is flagged.  Half a warp per voxel: lane = (th candidate,
// kw, tw candidate); the reachable planes (<= 3 per layer) are walked in a short loop.
struct MarkParams {
  int B, D, H, W, tiles_h, tiles_w, n_layers;
  int Din[3];
  uint16_t reach[3][kMaxPlanes];  // reach[l][p]: planes of layer l+1's input fed by plane p of layer l's
  uint8_t* flags[3];
};

__global__ void __launch_bounds__(256)
    k_mark_boxes(const long long* __restrict__ coords, int64_t K, const int32_t* __restrict__ n_voxels_dev,
                 const __grid_constant__ MarkParams Q) {
  int64_t total = K;
  if (n_voxels_dev != nullptr) {
    const int64_t live = *n_voxels_dev;
    if (live < total) total = live;
  }
  const uint32_t l16 = threadIdx.x & 15;
  const int th_sel = (l16 >> 3) & 1, kw = (l16 >> 1) & 3, cand = l16 & 1;
  const int64_t first = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 4;
  const int64_t step = ((int64_t)gridDim.x * blockDim.x) >> 4;
  const int plane_stride = Q.tiles_h * Q.tiles_w * 3;
  for (int64_t r = first; r < total; r += step) {
    const longlong2 c01 = *reinterpret_cast<const longlong2*>(coords + r * 4);
    const longlong2 c23 = *reinterpret_cast<const longlong2*>(coords + r * 4 + 2);
    const long long cb = c01.x, cz = c01.y, cy = c23.x, cx = c23.y;
    if (cb < 0 || cb >= Q.B || cz < 0 || cz >= Q.D || cy < 0 || cy >= Q.H || cx < 0 || cx >= Q.W || kw == 3) continue;
    const int b = (int)cb, y = (int)cy, x = (int)cx;
    uint32_t planes = 1u << (int)cz;
#pragma unroll
    for (int l = 0; l < 3; ++l) {
      if (l < Q.n_layers) {
        // rows [y-l, y+l] meet box rows [16 th - 1, 16 th + 16]
        int th_lo = (y - l - 1) >> 4, th_hi = (y + l + 1) >> 4;
        if (th_lo < 0) th_lo = 0;
        if (th_hi > Q.tiles_h - 1) th_hi = Q.tiles_h - 1;
        const int th = th_lo + th_sel;
        // columns [x-l, x+l] meet box columns [8 tw + kw - 1, 8 tw + kw + 6]:  8 tw in [s_lo, s_hi]
        const int s_lo = x - l - 6 - kw, s_hi = x + l + 1 - kw;
        int tw = (s_lo + 7) >> 3;  // ceil(s_lo / 8)
        if (tw < 0) tw = 0;
        tw += cand;
        const bool ok = th <= th_hi && tw * 8 <= s_hi && tw <= Q.tiles_w - 1;
        uint8_t* base = Q.flags[l] + (size_t)b * Q.Din[l] * plane_stride + (th * Q.tiles_w + tw) * 3 + kw;
        uint32_t pl = planes, next = 0;
        while (pl) {
          const int p = __ffs(pl) - 1;
          pl &= pl - 1;
          next |= Q.reach[l][p];
          if (ok) {
            uint8_t* f = base + (size_t)p * plane_stride;
            if (*f == 0) *f = 1;
          }
        }
        planes = next;
      }
    }
  }
}

// Per-plane schedule: which taps are valid, where their weights sit.
template <typename Params>
int plan_planes(Params& P, int stride_d, int pad_d) {
  const KdClasses kc = kd_classes(stride_d);
  bool fed[kSlots] = {false, false, false, false, false, false, false, false};
  for (int p = 0; p < P.Din; ++p) {
    PlaneInfo& pi = P.plane[p];
    pi = PlaneInfo{0, 0, 0, 0};
    const int cls = stride_d == 1 ? 0 : (((p + pad_d) % 2 == 0) ? 0 : 1);
    int jfirst = -1, jlast = -1, dmin = 0;
    for (int j = 0; j < kc.nkd[cls]; ++j) {  // kd descending -> d ascending
      const int num = p + pad_d - kc.kd[cls][j];
      if (num < 0 || num % stride_d != 0) continue;
      const int d = num / stride_d;
      if (d >= P.Dout) continue;
      if (jfirst < 0) { jfirst = j; dmin = d; }
      if (jlast >= 0 && j != jlast + 1) return -1;  // taps must be contiguous
      jlast = j;
      if (d != dmin + (j - jfirst)) return -1;      // and feed consecutive depths
      fed[d] = true;
    }
    if (jfirst < 0) continue;
    pi.nkd = jlast - jfirst + 1;
    pi.dmin = dmin;
    pi.woff = kc.base[cls] + (uint32_t)jfirst * kBTapBytes;
    pi.wstride = (uint32_t)kc.nkd[cls] * kBTapBytes;
  }
  for (int d = 0; d < P.Dout; ++d)
    if (!fed[d]) return -1;
  return 0;
}

// Transposed depth geometry (the data gradient of a forward layer with depth stride s, padding pad): plane d'
// of dY feeds dX planes p = d'*s - pad + kd, kd = 0..2 ascending - consecutive, so the taps are fused along N like
// the forward's.  Weights: the stride-1 image of W' = W.flip(kd,kh,kw).transpose(0,1) (tile j of the single kd
// class holds W'[kd' = 2-j] = W[kd = j]).
template <typename Params>
int plan_planes_transposed(Params& P, int stride_d, int pad_d) {
  for (int p = 0; p < kMaxPlanes; ++p) P.plane[p] = PlaneInfo{0, 0, 0, 0};
  for (int d = 0; d < P.Din; ++d) {
    int jfirst = -1, jlast = -1;
    for (int j = 0; j < 3; ++j) {
      const int p = d * stride_d - pad_d + j;
      if (p < 0 || p >= P.Dout) continue;
      if (jfirst < 0) jfirst = j;
      jlast = j;
    }
    if (jfirst < 0) continue;
    PlaneInfo& pi = P.plane[d];
    pi.nkd = jlast - jfirst + 1;
    pi.dmin = d * stride_d - pad_d + jfirst;
    pi.woff = (uint32_t)jfirst * kBTapBytes;
    pi.wstride = 3u * kBTapBytes;
  }
  return 0;
}

// 2-D convolution on the same kernel (RPN blocks, voxelnet/model.py:212-251): ONE input plane whose "depth taps" are the
// G groups of 64 output channels - one MMA of N = 64 G multiplies the box with the G weight tiles laid side by side and
// lands in G adjacent accumulators ("output depth" g = channels 64 g .. 64 g + 63).
template <typename Params>
int plan_planes_2d(Params& P, int groups) {
  for (int p = 0; p < kMaxPlanes; ++p) P.plane[p] = PlaneInfo{0, 0, 0, 0};
  if (groups < 1 || groups > 3) return -1;
  P.plane[0].nkd = groups;
  P.plane[0].dmin = 0;
  P.plane[0].woff = 0;
  P.plane[0].wstride = (uint32_t)groups * kBTapBytes;
  return 0;
}

// torch Conv2d weight [cout][cin][3][3] fp32 (times an optional per-output-channel scale: BatchNorm folded) -> the image
// the kernel reads for the schedule above: per (chunk, kw) a 72 KB slot holding [kh][group][n = 64][k = 64] bf16 tiles
// (rows 128B-swizzled); 64 floats of 1.0 behind the last slot serve as the epilogue's scale vector.
__global__ void __launch_bounds__(256)
    k_pack_weights2d(const float* __restrict__ w, const float* __restrict__ scale, int cin, int cout, int ksize,
                     uint8_t* __restrict__ packed) {
  const int nchunk = cin / 64, groups = cout / 64;
  const int total = ksize * ksize * cin * cout;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    int r = i;
    const int k = r % 64; r /= 64;
    const int n = r % 64; r /= 64;
    const int j = r % groups; r /= groups;
    const int kh = r % ksize; r /= ksize;
    const int kw = r % ksize; r /= ksize;
    const int c = r;
    const int co = j * 64 + n, ci = c * 64 + k;
    float v = w[(((size_t)co * cin + ci) * ksize + kh) * ksize + kw];
    if (scale != nullptr) v *= scale[co];
    const int kh3 = ksize == 1 ? 1 : kh, kw3 = ksize == 1 ? 1 : kw;  // a 1x1 kernel is the centre tap
    const size_t tile = (size_t)(c * 3 + kw3) * kBSlotBytes + (size_t)(kh3 * groups + j) * kBTapBytes;
    const size_t off = tile + (size_t)n * 128 + ((((k >> 3) ^ (n & 7)) << 4) | ((k & 7) << 1));
    *reinterpret_cast<__nv_bfloat16*>(packed + off) = __float2bfloat16_rn(v);
  }
  if (blockIdx.x == 0 && threadIdx.x < 64)
    reinterpret_cast<float*>(packed + (size_t)nchunk * 3 * kBSlotBytes)[threadIdx.x] = 1.0f;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

unsigned long long* g_debug_prof = nullptr;
int g_debug_flags = 0;
unsigned long long* g_debug_trace = nullptr;
int g_debug_trace_n = 0;

int out_depth(int Din, int stride_d, int pad_d) { return (Din + 2 * pad_d - 3) / stride_d + 1; }

struct PlanLayout {
  size_t groups_off, tiles_off, gtiles_off, total;
};
PlanLayout plan_layout(int64_t n_tiles) {
  PlanLayout L;
  L.groups_off = sizeof(PlanHeader);
  L.tiles_off = L.groups_off + (size_t)n_tiles * sizeof(GroupDesc);
  L.gtiles_off = L.tiles_off + (size_t)n_tiles * sizeof(TilePlan);
  L.total = L.gtiles_off + (size_t)n_tiles * kMaxGroupTiles * sizeof(TilePlan);  // (at most one group per tile)
  return L;
}

}  // namespace
}  // namespace vn

using namespace vn;

extern "C" {

size_t vn_conv3d_box_flags_bytes(int32_t B, int32_t Din, int32_t H, int32_t W) {
  if (B < 1 || Din < 1 || H < 1 || W < 1 || W % kTileW != 0) return 0;
  return (size_t)B * Din * ((H + kTileH - 1) / kTileH) * (W / kTileW) * 3;
}

size_t vn_conv3d_workspace_bytes(int32_t B, int32_t H, int32_t W) {
  if (B < 1 || H < 1 || W < 1 || W % kTileW != 0) return 0;
  return plan_layout((int64_t)B * ((H + kTileH - 1) / kTileH) * (W / kTileW)).total;
}

size_t vn_conv3d_out_state_bytes(int32_t B, int32_t H, int32_t W) {
  if (B < 1 || H < 1 || W < 1 || W % kTileW != 0) return 0;
  return (size_t)B * ((H + kTileH - 1) / kTileH) * (W / kTileW);
}

/* Debug aid (not in the public header's contract): the next conv launches write per-CTA cycle counters
 * into buf [148][16] u64: producer {total, wait empty_a, wait empty_b, boxes}, MMA issuer {total, wait
 * full_a, wait full_b, wait acc_empty}, epilogue warp 2 {total, wait sched, wait acc_full}.  NULL: off. */
void vn_conv3d_debug_profile(void* buf) { g_debug_prof = static_cast<unsigned long long*>(buf); }
/* Debug aid: experiment switches of the instrumented instantiation (1: no input-box loads, 2: no weight loads, 4: no
 * MMAs, 8: no epilogue stores).  Results are wrong while any is set; a non-NULL profile buffer or a non-zero flag word
 * selects the instrumented kernel, otherwise the production instantiation (no counters) runs. */
void vn_conv3d_debug_flags(int32_t flags) { g_debug_flags = flags; }
/* Debug aid: per-message timestamps of CTA 0 of the next conv launches into buf [n][4] u64 (producer acquired the stage,
 * issuer saw it, issuer done with it, message flags).  Only read by the instrumented instantiation. */
void vn_conv3d_debug_trace(void* buf, int32_t n) { g_debug_trace = static_cast<unsigned long long*>(buf); g_debug_trace_n = n; }

size_t vn_conv3d_packed_weight_bytes(int32_t cin) {
  if (cin != 64 && cin != 128) return 0;
  return (size_t)27 * cin * 64 * 2;
}

int vn_conv3d_pack_weights(const float* weight, int32_t cin, int32_t stride_d, void* packed, vn_stream_t stream) {
  VN_REQUIRE(cin == 64 || cin == 128, "conv3d_pack_weights: cin=%d (supported: 64, 128)", cin);
  VN_REQUIRE(stride_d == 1 || stride_d == 2, "conv3d_pack_weights: stride_d=%d (supported: 1, 2)", stride_d);
  VN_REQUIRE(weight != nullptr && packed != nullptr, "conv3d_pack_weights: null pointer");
  VN_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 15) == 0, "conv3d_pack_weights: packed must be 16-byte aligned");
  k_pack_weights<<<sm_count() * 4, 256, 0, as_stream(stream)>>>(weight, cin, stride_d, reinterpret_cast<uint8_t*>(packed));
  VN_LAUNCH_CHECK("k_pack_weights");
  return VN_OK;
}

int vn_conv3d_fold_background(const float* weight, int32_t cin, int32_t Din, int32_t stride_d, int32_t pad_d,
                              const float* scale, const float* shift, const float* bg_in, int32_t bg_classes,
                              float* shift_table, float* bg_out, vn_stream_t stream) {
  VN_REQUIRE(weight && scale && shift && shift_table, "conv3d_fold_background: null pointer");
  VN_REQUIRE(cin == 64 || cin == 128, "conv3d_fold_background: cin=%d (supported: 64, 128)", cin);
  VN_REQUIRE((stride_d == 1 || stride_d == 2) && (pad_d == 0 || pad_d == 1), "conv3d_fold_background: stride_d=%d pad_d=%d",
             stride_d, pad_d);
  VN_REQUIRE(Din >= 1 && Din <= kMaxPlanes && Din + 2 * pad_d >= 3, "conv3d_fold_background: Din=%d", Din);
  VN_REQUIRE(bg_classes == 1 || (bg_classes == 9 && bg_in != nullptr && bg_out == nullptr),
             "conv3d_fold_background: bg_classes=%d (1, or 9 with bg_in and without bg_out)", bg_classes);
  const int Dout = out_depth(Din, stride_d, pad_d);
  VN_REQUIRE(Dout >= 1 && Dout <= kSlots, "conv3d_fold_background: Dout=%d outside [1,%d]", Dout, kSlots);
  const int nv = bg_classes == 9 ? 25 : 9;
  k_fold_background<<<nv * Dout, 64, 0, as_stream(stream)>>>(weight, cin, Din, Dout, stride_d, pad_d, scale, shift, bg_in,
                                                             bg_classes, shift_table, bg_out);
  VN_LAUNCH_CHECK("k_fold_background");
  return VN_OK;
}

int vn_conv3d_mark_boxes(const int64_t* coords, int64_t K, const int32_t* n_voxels_dev, int32_t B, int32_t D, int32_t H,
                         int32_t W, int32_t n_layers, const int32_t* stride_d, const int32_t* pad_d,
                         uint8_t* const* flags, vn_stream_t stream) {
  VN_REQUIRE(K >= 0, "conv3d_mark_boxes: K < 0");
  VN_REQUIRE(n_layers >= 1 && n_layers <= 3 && stride_d && pad_d && flags, "conv3d_mark_boxes: n_layers=%d", n_layers);
  VN_REQUIRE(B >= 1 && D >= 1 && D <= kMaxPlanes && H >= 1 && W >= 1 && W % kTileW == 0, "conv3d_mark_boxes: bad grid");
  if (K == 0) return VN_OK;
  VN_REQUIRE(coords != nullptr, "conv3d_mark_boxes: coords is null");
  MarkParams Q{};
  Q.B = B; Q.D = D; Q.H = H; Q.W = W;
  Q.tiles_h = (H + kTileH - 1) / kTileH;
  Q.tiles_w = W / kTileW;
  Q.n_layers = n_layers;
  int din = D;
  for (int l = 0; l < n_layers; ++l) {
    VN_REQUIRE((stride_d[l] == 1 || stride_d[l] == 2) && (pad_d[l] == 0 || pad_d[l] == 1) && din + 2 * pad_d[l] >= 3,
               "conv3d_mark_boxes: layer %d geometry", l);
    VN_REQUIRE(flags[l] != nullptr, "conv3d_mark_boxes: flags[%d] is null", l);
    Q.Din[l] = din; Q.flags[l] = flags[l];
    const int dout = out_depth(din, stride_d[l], pad_d[l]);
    VN_REQUIRE(dout >= 1 && dout <= kMaxPlanes, "conv3d_mark_boxes: layer %d output depth %d", l, dout);
    for (int p = 0; p < din; ++p) {
      uint16_t m = 0;
      for (int d = 0; d < dout; ++d)
        for (int kd = 0; kd < 3; ++kd)
          if (d * stride_d[l] - pad_d[l] + kd == p) m |= (uint16_t)(1u << d);
      Q.reach[l][p] = m;
    }
    din = dout;
  }
  const int64_t want = (K + 15) / 16;  // 16 voxels per 256-thread block
  const int64_t cap = (int64_t)sm_count() * 8;
  k_mark_boxes<<<(unsigned)(want < cap ? want : cap), 256, 0, as_stream(stream)>>>(
      reinterpret_cast<const long long*>(coords), K, n_voxels_dev, Q);
  VN_LAUNCH_CHECK("k_mark_boxes");
  return VN_OK;
}

// Input of the gather instantiation (vn_conv3d_bn_relu_fwd_rows): occupancy index + compact row table instead of x.
struct GatherSrc {
  const uint32_t* bitmap;
  const uint32_t* prefix;
  size_t words;
  const int32_t* voxel_offsets;
  const void* rows;
  int64_t max_voxels;
};

static int conv_launch(const void* x, int32_t B, int32_t Din, int32_t H, int32_t W, int32_t cin, int32_t stride_d,
                       int32_t pad_d, const void* packed_weight, const float* scale, const float* shift_table,
                       int32_t n_variants, const float* sub, int32_t out_mode, void* y, uint8_t* box_flags,
                       int32_t flag_opts, uint8_t* out_state, void* workspace, size_t workspace_bytes,
                       vn_stream_t stream, double* stats, int32_t transposed_dout = 0, const GatherSrc* gather = nullptr,
                       int32_t groups2d = 0, int32_t out_ctotal = 0, int32_t out_coff = 0, int32_t s2_hin = 0, int32_t s2_win = 0,
                       int32_t pointwise = 0, int32_t out_up = 1, int32_t out_oy = 0, int32_t out_ox = 0, int32_t no_relu = 0) {
  // s2_hin > 0 (2-D layers only): stride 2 in H and W; x is [B, 1, s2_hin, s2_win, cin] and H, W are the OUTPUT extents
  VN_REQUIRE((x || gather) && packed_weight && scale && shift_table && y, "conv3d: null pointer");
  VN_REQUIRE(cin == 64 || cin == 128 || (groups2d > 0 && cin % 64 == 0 && cin >= 64 && cin <= 1024),
             "conv3d: cin=%d (supported: 64, 128; 2-D layers: multiples of 64 up to 1024)", cin);
  VN_REQUIRE(B >= 1 && Din >= 1 && H >= 1 && W >= 1, "conv3d: empty input");
  VN_REQUIRE((stride_d == 1 || stride_d == 2) && (pad_d == 0 || pad_d == 1), "conv3d: stride_d=%d pad_d=%d unsupported",
             stride_d, pad_d);
  VN_REQUIRE(out_mode == 0 || out_mode == 1 || out_mode == 2, "conv3d: out_mode=%d", out_mode);
  VN_REQUIRE((W % kTileW == 0 || groups2d > 0) && (W + kTileW - 1) / kTileW <= kMaxTilesW,
             "conv3d: W=%d must be a multiple of %d (3-D layers), at most %d", W, kTileW, kTileW * kMaxTilesW);
  VN_REQUIRE(Din <= kMaxPlanes, "conv3d: Din=%d exceeds %d", Din, kMaxPlanes);
  VN_REQUIRE(groups2d == 0 || (Din == 1 && groups2d >= 1 && groups2d <= 3), "conv2d: %d output channel groups (1..3), Din=%d", groups2d, Din);
  const int Dout = groups2d > 0 ? groups2d : (transposed_dout > 0 ? transposed_dout : out_depth(Din, stride_d, pad_d));
  VN_REQUIRE((transposed_dout > 0 || groups2d > 0 || Din + 2 * pad_d >= 3) && Dout >= 1 && Dout <= kSlots, "conv3d: Dout=%d outside [1,%d]",
             Dout, kSlots);
  VN_REQUIRE(n_variants == 0 || n_variants == 1 || n_variants == 9 || n_variants == 25,
             "conv3d: n_variants=%d (0: plain [64] shift, 1, 9, 25)", n_variants);
  const int nvar = n_variants == 9 ? 9 : (n_variants == 25 ? 25 : 1);
  VN_REQUIRE(nvar == 1 || (H >= 5 && W >= 5), "conv3d: border variants need H, W >= 5");
  VN_REQUIRE((reinterpret_cast<uintptr_t>(gather ? gather->rows : x) & 15) == 0 && (reinterpret_cast<uintptr_t>(packed_weight) & 15) == 0 &&
                 (reinterpret_cast<uintptr_t>(y) & 15) == 0,
             "conv3d: buffers must be 16-byte aligned");
  VN_REQUIRE(workspace != nullptr && (reinterpret_cast<uintptr_t>(workspace) & 15) == 0, "conv3d: workspace missing or misaligned");

  const int tiles_h = (H + kTileH - 1) / kTileH, tiles_w = (W + kTileW - 1) / kTileW;
  const int64_t n_tiles = (int64_t)B * tiles_h * tiles_w;
  VN_REQUIRE(n_tiles < (int64_t)0x7fffffff, "conv3d: too many tiles");
  const PlanLayout L = plan_layout(n_tiles);
  if (workspace_bytes < L.total) {
    set_error("conv3d: workspace %zu < %zu bytes", workspace_bytes, L.total);
    return VN_ERR_WORKSPACE;
  }
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  PlanHeader* hdr = reinterpret_cast<PlanHeader*>(ws);
  GroupDesc* groups = reinterpret_cast<GroupDesc*>(ws + L.groups_off);
  TilePlan* tiles = reinterpret_cast<TilePlan*>(ws + L.tiles_off);
  TilePlan* gtiles = reinterpret_cast<TilePlan*>(ws + L.gtiles_off);

  CUtensorMap tmap;
  memset(&tmap, 0, sizeof(tmap));
  if (gather == nullptr) {
  EncodeTiledFn encode = encode_tiled_fn();
  if (encode == nullptr) {
    set_error("conv3d: cuTensorMapEncodeTiled not available from the driver");
    return VN_ERR_DRIVER;
  }
  const bool s2 = s2_hin > 0;
  VN_REQUIRE(!s2 || (groups2d > 0 && gather == nullptr && stats == nullptr && (s2_hin - 1) / 2 + 1 == H && (s2_win - 1) / 2 + 1 == W),
             "conv2d: stride-2 input %d x %d does not give a %d x %d output", s2_hin, s2_win, H, W);
  const cuuint64_t Hx = s2 ? s2_hin : H, Wx = s2 ? s2_win : W;
  const cuuint64_t gdim[5] = {(cuuint64_t)cin, Wx, Hx, (cuuint64_t)Din, (cuuint64_t)B};
  const cuuint64_t gstr[4] = {(cuuint64_t)cin * 2, Wx * cin * 2, Hx * Wx * cin * 2, (cuuint64_t)Din * Hx * Wx * cin * 2};
  // stride 2: the box spans 16 columns x 34 rows of the input and loads every second one (8 x 17 samples)
  const cuuint32_t box[5] = {64, (cuuint32_t)(s2 ? 2 * kTileW : kTileW), (cuuint32_t)(s2 ? 2 * (kTileH + 1) : kHaloH), 1, 1};
  const cuuint32_t estr[5] = {1, (cuuint32_t)(s2 ? 2 : 1), (cuuint32_t)(s2 ? 2 : 1), 1, 1};
  const CUresult cr = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(x), gdim, gstr, box, estr,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                             CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (cr != CUDA_SUCCESS) {
    set_error("conv3d: cuTensorMapEncodeTiled failed with CUresult %d", (int)cr);
    return VN_ERR_DRIVER;
  }
  }

  cudaStream_t st = as_stream(stream);
  PlanParams Q{};
  Q.B = B; Q.Din = Din; Q.Dout = Dout; Q.H = H; Q.W = W; Q.tiles_h = tiles_h; Q.tiles_w = tiles_w;
  Q.box_flags = box_flags;
  Q.consume = (flag_opts & VN_CONV_FLAGS_CONSUME) ? 1 : 0;
  Q.force_ring = (flag_opts & VN_CONV_FLAGS_BORDER_RING) ? 1 : 0;
  Q.out_state = out_state;
  Q.only_kw1 = pointwise ? 1 : 0;
  Q.hdr = hdr; Q.groups = groups; Q.tiles = tiles; Q.gtiles = gtiles;
  if ((groups2d > 0 ? plan_planes_2d(Q, groups2d)
                    : (transposed_dout > 0 ? plan_planes_transposed(Q, stride_d, pad_d) : plan_planes(Q, stride_d, pad_d))) != 0) {
    set_error("conv3d: unsupported depth schedule (Din=%d stride_d=%d pad_d=%d)", Din, stride_d, pad_d);
    return VN_ERR_UNSUPPORTED;
  }
  if ((flag_opts & VN_CONV_FLAGS_PLANNED) && groups2d > 0) {
    // a dense 2-D plan depends on (B, H, W, groups) only: reuse it, rewind the work counter
    VN_CUDA(cudaMemsetAsync(&hdr->next_group, 0, sizeof(int32_t), st));
  }
  if (!(flag_opts & VN_CONV_FLAGS_PLANNED)) {  // else: vn_conv3d_plan already built the plan in this workspace
    VN_CUDA(cudaMemsetAsync(hdr, 0, sizeof(PlanHeader), st));
    k_conv3d_plan<<<B * tiles_h, 64, 0, st>>>(Q);
    VN_LAUNCH_CHECK("k_conv3d_plan");
  }

  ConvParams P{};
  P.B = B; P.Din = Din; P.Dout = Dout; P.H = H; P.W = W;
  P.nchunk = cin / 64;
  P.tiles_h = tiles_h; P.tiles_w = tiles_w; P.n_tiles = (int)n_tiles;
  P.out_mode = out_mode;
  P.nvar = nvar;
  P.table_smem = nvar * Dout <= kMaxTableRows ? 1 : 0;
  P.sub_classes = (nvar == 9 && sub != nullptr) ? 9 : 1;
  P.wpack = reinterpret_cast<const uint8_t*>(packed_weight);
  P.scale = scale;
  P.table = shift_table;
  P.table_plain = n_variants == 0 ? 1 : 0;
  P.reverse = (flag_opts & VN_CONV_FLAGS_REVERSE) ? 1 : 0;
  P.out_ctotal = out_ctotal > 0 ? out_ctotal : 64 * Dout;
  P.out_coff = out_coff;
  P.out_up = out_up; P.out_oy = out_oy; P.out_ox = out_ox;
  P.pointwise = pointwise;
  P.relu_floor = no_relu ? -INFINITY : 0.0f;
  P.sub = sub;
  P.out = y;
  P.hdr = hdr; P.groups = groups; P.tiles = tiles; P.gtiles = gtiles;
  P.prof = g_debug_prof;
  P.debug = g_debug_flags;
  P.trace = g_debug_trace;
  P.trace_n = g_debug_trace_n;
  for (int p = 0; p < kMaxPlanes; ++p) P.plane[p] = Q.plane[p];

  const int grid = n_tiles < sm_count() ? (int)n_tiles : sm_count();
  P.stats = stats;
  if (gather != nullptr) {
    VN_REQUIRE(stats == nullptr && cin == 128, "conv3d: the row-table input serves the first inference layer only");
    VN_REQUIRE(nvar * Dout <= kGatherTableRows, "conv3d: the row-table input takes a plain or per-depth shift (%d x %d table rows > %d)", nvar, Dout, kGatherTableRows);
    P.g_bitmap = gather->bitmap; P.g_prefix = gather->prefix; P.g_words = (uint32_t)gather->words;
    P.g_voxoff = gather->voxel_offsets;
    P.g_rows = reinterpret_cast<const __nv_bfloat16*>(gather->rows);
    P.g_max_voxels = gather->max_voxels;
    if (P.prof != nullptr || P.debug != 0) {
      VN_CUDA(cudaFuncSetAttribute(k_conv3d_bn_relu<false, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
      k_conv3d_bn_relu<false, true, true><<<grid, kThreadsGather, kSmemBytes, st>>>(tmap, P);
    } else {
      VN_CUDA(cudaFuncSetAttribute(k_conv3d_bn_relu<false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
      k_conv3d_bn_relu<false, false, true><<<grid, kThreadsGather, kSmemBytes, st>>>(tmap, P);
    }
  } else if (s2_hin > 0 || pointwise) {
    VN_CUDA(cudaFuncSetAttribute(k_conv3d_bn_relu<false, false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    k_conv3d_bn_relu<false, false, false, true><<<grid, kThreads, kSmemBytes, st>>>(tmap, P);
  } else if (stats != nullptr) {
    VN_CUDA(cudaFuncSetAttribute(k_conv3d_bn_relu<true, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    k_conv3d_bn_relu<true, false, false><<<grid, kThreads, kSmemBytes, st>>>(tmap, P);
  } else if (P.prof != nullptr || P.debug != 0) {
    VN_CUDA(cudaFuncSetAttribute(k_conv3d_bn_relu<false, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    k_conv3d_bn_relu<false, true, false><<<grid, kThreads, kSmemBytes, st>>>(tmap, P);
  } else {
    VN_CUDA(cudaFuncSetAttribute(k_conv3d_bn_relu<false, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    k_conv3d_bn_relu<false, false, false><<<grid, kThreads, kSmemBytes, st>>>(tmap, P);
  }
  VN_LAUNCH_CHECK("k_conv3d_bn_relu");
  return VN_OK;
}

int vn_conv3d_bn_relu_fwd(const void* x, int32_t B, int32_t Din, int32_t H, int32_t W, int32_t cin, int32_t stride_d,
                          int32_t pad_d, const void* packed_weight, const float* scale, const float* shift_table,
                          int32_t n_variants, const float* sub, int32_t out_mode, void* y, uint8_t* box_flags,
                          int32_t flag_opts, uint8_t* out_state, void* workspace, size_t workspace_bytes,
                          vn_stream_t stream) {
  return conv_launch(x, B, Din, H, W, cin, stride_d, pad_d, packed_weight, scale, shift_table, n_variants, sub, out_mode, y,
                     box_flags, flag_opts, out_state, workspace, workspace_bytes, stream, nullptr);
}

int vn_conv3d_bn_relu_fwd_rows(const vn_grid_t* grid, const void* rows_bf16, int64_t max_voxels, const int32_t* voxel_offsets,
                               int32_t n_frames, int64_t n_points_total, const void* vox_workspace,
                               size_t vox_workspace_bytes, int32_t stride_d, int32_t pad_d, const void* packed_weight,
                               const float* scale, const float* shift_table, int32_t n_variants, const float* sub,
                               int32_t out_mode, void* y, uint8_t* box_flags, int32_t flag_opts, uint8_t* out_state,
                               void* workspace, size_t workspace_bytes, vn_stream_t stream) {
  VN_REQUIRE(grid != nullptr && rows_bf16 != nullptr && voxel_offsets != nullptr && vox_workspace != nullptr,
             "conv3d_rows: null pointer");
  VN_REQUIRE(max_voxels >= 0 && n_frames >= 1 && n_points_total >= 0, "conv3d_rows: bad sizes");
  VN_REQUIRE(grid->D >= 1 && grid->D <= kMaxPlanes && (int64_t)grid->D * grid->H * grid->W < ((int64_t)1 << 31),
             "conv3d_rows: grid depth %d outside [1,%d] or too many cells", grid->D, kMaxPlanes);
  GatherSrc g{};
  if (occupancy_views(*grid, n_frames, n_points_total, vox_workspace, vox_workspace_bytes, &g.bitmap, &g.prefix, &g.words) != 0) {
    set_error("conv3d_rows: voxelizer workspace of %zu bytes is too small for this batch", vox_workspace_bytes);
    return VN_ERR_WORKSPACE;
  }
  g.voxel_offsets = voxel_offsets;
  g.rows = rows_bf16;
  g.max_voxels = max_voxels;
  return conv_launch(nullptr, n_frames, grid->D, grid->H, grid->W, 128, stride_d, pad_d, packed_weight, scale, shift_table,
                     n_variants, sub, out_mode, y, box_flags, flag_opts, out_state, workspace, workspace_bytes, stream, nullptr,
                     0, &g);
}

int vn_conv3d_plan(int32_t B, int32_t Din, int32_t H, int32_t W, int32_t stride_d, int32_t pad_d, uint8_t* box_flags,
                   int32_t flag_opts, uint8_t* out_state, void* workspace, size_t workspace_bytes, vn_stream_t stream) {
  VN_REQUIRE(B >= 1 && Din >= 1 && Din <= kMaxPlanes && H >= 1 && W >= 1, "conv3d_plan: bad shape");
  VN_REQUIRE((stride_d == 1 || stride_d == 2) && (pad_d == 0 || pad_d == 1), "conv3d_plan: stride_d=%d pad_d=%d unsupported",
             stride_d, pad_d);
  VN_REQUIRE(W % kTileW == 0 && W / kTileW <= kMaxTilesW, "conv3d_plan: W=%d must be a multiple of %d, at most %d", W, kTileW,
             kTileW * kMaxTilesW);
  const int Dout = out_depth(Din, stride_d, pad_d);
  VN_REQUIRE(Din + 2 * pad_d >= 3 && Dout >= 1 && Dout <= kSlots, "conv3d_plan: Dout=%d outside [1,%d]", Dout, kSlots);
  VN_REQUIRE(workspace != nullptr && (reinterpret_cast<uintptr_t>(workspace) & 15) == 0, "conv3d_plan: workspace missing or misaligned");
  const int tiles_h = (H + kTileH - 1) / kTileH, tiles_w = W / kTileW;
  const int64_t n_tiles = (int64_t)B * tiles_h * tiles_w;
  VN_REQUIRE(n_tiles < (int64_t)0x7fffffff, "conv3d_plan: too many tiles");
  const PlanLayout L = plan_layout(n_tiles);
  if (workspace_bytes < L.total) {
    set_error("conv3d_plan: workspace %zu < %zu bytes", workspace_bytes, L.total);
    return VN_ERR_WORKSPACE;
  }
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  PlanParams Q{};
  Q.B = B; Q.Din = Din; Q.Dout = Dout; Q.H = H; Q.W = W; Q.tiles_h = tiles_h; Q.tiles_w = tiles_w;
  Q.box_flags = box_flags;
  Q.consume = (flag_opts & VN_CONV_FLAGS_CONSUME) ? 1 : 0;
  Q.force_ring = (flag_opts & VN_CONV_FLAGS_BORDER_RING) ? 1 : 0;
  Q.out_state = out_state;
  Q.hdr = reinterpret_cast<PlanHeader*>(ws);
  Q.groups = reinterpret_cast<GroupDesc*>(ws + L.groups_off);
  Q.tiles = reinterpret_cast<TilePlan*>(ws + L.tiles_off);
  Q.gtiles = reinterpret_cast<TilePlan*>(ws + L.gtiles_off);
  if (plan_planes(Q, stride_d, pad_d) != 0) {
    set_error("conv3d_plan: unsupported depth schedule (Din=%d stride_d=%d pad_d=%d)", Din, stride_d, pad_d);
    return VN_ERR_UNSUPPORTED;
  }
  cudaStream_t st = as_stream(stream);
  VN_CUDA(cudaMemsetAsync(Q.hdr, 0, sizeof(PlanHeader), st));
  k_conv3d_plan<<<B * tiles_h, 64, 0, st>>>(Q);
  VN_LAUNCH_CHECK("k_conv3d_plan");
  return VN_OK;
}

int vn_conv3d_dgrad(const void* dy, int32_t B, int32_t Dout_fwd, int32_t H, int32_t W, int32_t Din_fwd, int32_t stride_d,
                    int32_t pad_d, const void* packed_weight_t, const float* ones, const float* zeros, void* dx,
                    double* stats_scratch, void* workspace, size_t workspace_bytes, vn_stream_t stream) {
  VN_REQUIRE(stats_scratch && ones && zeros, "conv3d_dgrad: null pointer");
  VN_REQUIRE(Din_fwd >= 1 && Din_fwd <= kSlots, "conv3d_dgrad: Din_fwd=%d outside [1,%d]", Din_fwd, kSlots);
  return conv_launch(dy, B, Dout_fwd, H, W, 64, stride_d, pad_d, packed_weight_t, ones, zeros, 0, nullptr, 0, dx, nullptr, 0,
                     nullptr, workspace, workspace_bytes, stream, stats_scratch, Din_fwd);
}

int vn_conv3d_raw_stats_fwd(const void* x, int32_t B, int32_t Din, int32_t H, int32_t W, int32_t cin, int32_t stride_d,
                            int32_t pad_d, const void* packed_weight, const float* ones, const float* bias, void* y_raw,
                            double* stats, uint8_t* box_flags, int32_t flag_opts, void* workspace,
                            size_t workspace_bytes, vn_stream_t stream) {
  VN_REQUIRE(stats != nullptr && ones != nullptr && bias != nullptr, "conv3d_raw_stats: null pointer");
  VN_REQUIRE((reinterpret_cast<uintptr_t>(stats) & 7) == 0, "conv3d_raw_stats: stats must be 8-byte aligned");
  return conv_launch(x, B, Din, H, W, cin, stride_d, pad_d, packed_weight, ones, bias, 0, nullptr, 0, y_raw, box_flags,
                     flag_opts, nullptr, workspace, workspace_bytes, stream, stats);
}

size_t vn_conv2d_workspace_bytes(int32_t B, int32_t H, int32_t W) {
  if (B < 1 || H < 1 || W < 1) return 0;
  return plan_layout((int64_t)B * ((H + kTileH - 1) / kTileH) * ((W + kTileW - 1) / kTileW)).total;
}

size_t vn_conv2d_packed_weight_bytes(int32_t cin) {
  if (cin < 64 || cin > 1024 || cin % 64 != 0) return 0;
  return (size_t)(cin / 64) * 3 * kBSlotBytes + 64 * sizeof(float);
}

int vn_conv2d_pack_weights(const float* weight, const float* scale, int32_t cin, int32_t cout, int32_t ksize, void* packed,
                           vn_stream_t stream) {
  VN_REQUIRE(ksize == 3 || ksize == 1, "conv2d_pack_weights: ksize=%d (supported: 3, 1)", ksize);
  VN_REQUIRE(cin >= 64 && cin <= 1024 && cin % 64 == 0, "conv2d_pack_weights: cin=%d (multiples of 64 up to 1024)", cin);
  VN_REQUIRE(cout == 64 || cout == 128 || cout == 192, "conv2d_pack_weights: cout=%d (supported: 64, 128, 192)", cout);
  VN_REQUIRE(weight != nullptr && packed != nullptr && (reinterpret_cast<uintptr_t>(packed) & 15) == 0,
             "conv2d_pack_weights: null or misaligned pointer");
  VN_CUDA(cudaMemsetAsync(packed, 0, vn_conv2d_packed_weight_bytes(cin), as_stream(stream)));
  k_pack_weights2d<<<sm_count() * 2, 256, 0, as_stream(stream)>>>(weight, scale, cin, cout, ksize, reinterpret_cast<uint8_t*>(packed));
  VN_LAUNCH_CHECK("k_pack_weights2d");
  return VN_OK;
}

int vn_conv2d_bn_relu_fwd(const void* x, int32_t B, int32_t H, int32_t W, int32_t cin, int32_t cout, int32_t stride,
                          const void* packed_weight, const float* shift, void* y, int32_t y_channels, int32_t y_channel_offset,
                          int32_t planned, void* workspace, size_t workspace_bytes, vn_stream_t stream) {
  VN_REQUIRE(cout == 64 || cout == 128 || cout == 192, "conv2d: cout=%d (supported: 64, 128, 192)", cout);
  VN_REQUIRE(cin >= 64 && cin <= 1024 && cin % 64 == 0, "conv2d: cin=%d (multiples of 64 up to 1024)", cin);
  VN_REQUIRE(stride == 1 || stride == 2, "conv2d: stride=%d (supported: 1, 2)", stride);
  VN_REQUIRE(packed_weight != nullptr && shift != nullptr && H >= 1 && W >= 1, "conv2d: null pointer or empty input");
  VN_REQUIRE(y_channels >= cout && y_channel_offset >= 0 && y_channel_offset + cout <= y_channels && y_channels % 16 == 0 &&
                 y_channel_offset % 16 == 0,
             "conv2d: channels %d..%d of %d (multiples of 16 inside y)", y_channel_offset, y_channel_offset + cout, y_channels);
  const float* ones = reinterpret_cast<const float*>(static_cast<const uint8_t*>(packed_weight) + (size_t)(cin / 64) * 3 * kBSlotBytes);
  const int Ho = stride == 2 ? (H - 1) / 2 + 1 : H, Wo = stride == 2 ? (W - 1) / 2 + 1 : W;
  return conv_launch(x, B, 1, Ho, Wo, cin, 1, 1, packed_weight, ones, shift, 1, nullptr, 2, y, nullptr,
                     planned ? VN_CONV_FLAGS_PLANNED : 0, nullptr, workspace, workspace_bytes, stream, nullptr, 0, nullptr, cout / 64,
                     y_channels, y_channel_offset, stride == 2 ? H : 0, stride == 2 ? W : 0);
}

int vn_conv2d_pointwise_fwd(const void* x, int32_t B, int32_t H, int32_t W, int32_t cin, int32_t cout, const void* packed_weight,
                            const float* shift, int32_t relu, void* y, int32_t y_channels, int32_t y_channel_offset, int32_t up,
                            int32_t oy, int32_t ox, int32_t planned, void* workspace, size_t workspace_bytes, vn_stream_t stream) {
  VN_REQUIRE(cout == 64 || cout == 128 || cout == 192, "conv2d_pointwise: cout=%d (supported: 64, 128, 192)", cout);
  VN_REQUIRE(cin >= 64 && cin <= 1024 && cin % 64 == 0, "conv2d_pointwise: cin=%d (multiples of 64 up to 1024)", cin);
  VN_REQUIRE(packed_weight != nullptr && shift != nullptr && H >= 1 && W >= 1, "conv2d_pointwise: null pointer or empty input");
  VN_REQUIRE(up >= 1 && up <= 8 && oy >= 0 && oy < up && ox >= 0 && ox < up, "conv2d_pointwise: up=%d offset (%d,%d)", up, oy, ox);
  VN_REQUIRE(y_channels >= cout && y_channel_offset >= 0 && y_channel_offset + cout <= y_channels && y_channels % 16 == 0 &&
                 y_channel_offset % 16 == 0,
             "conv2d_pointwise: channels %d..%d of %d (multiples of 16 inside y)", y_channel_offset, y_channel_offset + cout, y_channels);
  const float* ones = reinterpret_cast<const float*>(static_cast<const uint8_t*>(packed_weight) + (size_t)(cin / 64) * 3 * kBSlotBytes);
  return conv_launch(x, B, 1, H, W, cin, 1, 1, packed_weight, ones, shift, 1, nullptr, 2, y, nullptr,
                     planned ? VN_CONV_FLAGS_PLANNED : 0, nullptr, workspace, workspace_bytes, stream, nullptr, 0, nullptr, cout / 64,
                     y_channels, y_channel_offset, 0, 0, 1, up, oy, ox, relu ? 0 : 1);
}

}  // extern "C"
