// Segment (CSR) aggregation kernels: the bf16 GCN / GraphSAGE / GraphConv message-passing hot loop.
//
// Replaces: P2 `zeros -> scatter_add(0, src.expand, m)` (experiments/OGB/GCN.py:57-65), the
// `x[src]`, `x[dst]`, `cat`, ReLU chain of P1 (GCN.py:43-52) after the algebraic split
//      relu(W [x_i ; x_j] + b) = relu(P[i] + Q[j]),  P = X W_i^T + b,  Q = X W_j^T,
// and the local halves of the reference's K6/K7 scatter/gather kernels (local_data_kernels.cuh:301-406).
//
// Design: one sub-warp group (G lanes) owns one output row and walks its CSR segment; every lane keeps
// 8 bf16 columns (one 16-byte load) in fp32 registers, so the reduction is atomic-free and deterministic.
// Neighbour indices are loaded coalesced and broadcast with shuffles; UNROLL independent 16-byte row loads
// are in flight per lane.  When HALO is set, a negative neighbour index names a row of ANOTHER rank's matrix
// (~idx = owner << peer_shift | row): the lane group reads it *directly from the owner GPU* through the NVLink peer
// pointer table (pull), in the same batch of loads as its local neighbours, so no halo buffer is materialised and
// the NVLink latency overlaps the local gathers.
#include <stdlib.h>

#include "common.cuh"

namespace dgb {

enum SegMode { SEG_SUM = 0, SEG_ADD_RELU = 1, SEG_ADD_RELU_BWD = 2, SEG_MASKED_SUM = 3 };

struct SegParams {
  // CSR (or CSC) structure: segment r covers slots [ptr[r], ptr_end[r]) of idx (ptr_end = ptr + 1 for a whole segment; other
  // arrays let a launch walk only a sub-range of every segment, e.g. the local-neighbour or the halo-neighbour part)
  const int* ptr;
  const int* ptr_end;
  // SEG_ADD_RELU: add to the existing out / cnt rows instead of overwriting them (second pass over the halo neighbours);
  // rows whose range is empty are left untouched
  int accumulate;
  const int* idx;
  int64_t n_rows;
  int H;  // columns (bf16), multiple of 8
  // gathered operand (rows indexed by idx); byte strides fit in 32 bits (checked on the host)
  const __nv_bfloat16* src;
  uint32_t ld_src_b;
  // second gathered operand (SEG_ADD_RELU_BWD: P rows indexed by idx)
  const __nv_bfloat16* src2;
  uint32_t ld_src2_b;
  // per-output-row operand (SEG_ADD_RELU: P[r]; SEG_ADD_RELU_BWD: Q[r])
  const __nv_bfloat16* rowop;
  int64_t ld_rowop;
  // optional per-slot weight (fp32) and mean normalisation
  const float* slot_w;
  int mean;
  // outputs
  __nv_bfloat16* out;
  int64_t ld_out;
  unsigned short* cnt;  // SEG_ADD_RELU: number of active (pre-activation > 0) edges per (row, col); may be null
  int64_t ld_cnt;
  // ReLU bit masks, one byte per (edge slot, 8-column chunk): written by SEG_ADD_RELU (indexed by this walk's slot),
  // read by SEG_MASKED_SUM through slot_map (slot of the *forward* walk for each slot of this walk)
  unsigned char* mask;
  const int* slot_map;
  // SEG_MASKED_SUM only: rows r < dp_rows additionally produce dP[r,:] = src[r,:] * cnt[r,:] (the gradient of the per-row
  // operand of the forward kernel) in the same pass, so no separate elementwise kernel re-reads dOut
  __nv_bfloat16* dp_out;
  int64_t ld_dp;
  int64_t dp_rows;
  // SEG_MASKED_SUM / SEG_SUM: output rows r >= out_split go to out2[(r - out_split), :] (the CSC walk of a distributed layer writes
  // the gradients of halo rows straight into the symmetric slot the owners reduce from; no copy)
  __nv_bfloat16* out2;
  int64_t ld_out2;
  int64_t out_split;
  int out2_skip_empty;  // rows >= out_split with an empty segment are not written (their reader knows they are structurally zero)
  // halo pull (HALO kernels): idx[k] < 0 encodes a remote row, ~idx[k] = (owner << peer_shift) | row in the owner's matrix
  const __nv_bfloat16* const* peer_base;  // [world] base pointer of the same matrix on every rank
  int world;
  int peer_shift;
  uint32_t ld_peer_b;
  int rows_per_group;  // consecutive rows per lane group (set by launch_seg)
};

constexpr int kMaxPeers = 64;

template <bool HALO>
__device__ __forceinline__ const char* src_row(const SegParams& p, const char* const* s_peer, const char* src_lane,
                                               uint32_t lane_off, int j) {
  if (HALO) {
    if (j < 0) {  // remote row: owner's base from the shared-memory copy of the peer table, no dependent global load
      const uint32_t c = ~(uint32_t)j;
      return s_peer[c >> p.peer_shift] + (uint64_t)(c & ((1u << p.peer_shift) - 1u)) * p.ld_peer_b + lane_off;
    }
  }
  return src_lane + (uint64_t)(uint32_t)j * p.ld_src_b;  // one IMAD.WIDE.U32
}

// HALO kernels issue ONE load form for local and remote rows (coherent, no L1 allocation, evict_last hint for the local L2; a peer
// address bypasses the local L2 anyway), so the batch of UNROLL loads stays branch-free.
__device__ __forceinline__ uint4 ldg_any_hint_v4(const void* p, uint64_t pol) {
  uint4 r;
  asm volatile("ld.global.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p), "l"(pol));
  return r;
}

// SEG_ADD_RELU on packed bf16 pairs (two columns per 32-bit lane word).  Per pair and edge: one HFMA2.BF16_V2.RELU forms the message
// m = relu(P[i] * 1 + Q[j]) (rounded to bf16 once -- exactly what a bf16 Linear -> ReLU produces; fma pipe), one HSET2 turns it into the
// activity mask (alu pipe), two FHADD.BF16 accumulate the halves in fp32 straight from the packed register (fma pipe; no shift / mask
// to widen them), one IADD counts and one LOP3 collects the mask bit.  History (ncu, 61.9 M edges x 256): scalar fp32 version 72 SASS
// instructions per edge-lane and issue-bound at 69 % SM throughput; HADD2 + HMNMX2 + HSET2 + shift / mask widening 47 (alu pipe 65 %);
// this form 35, three per pair on each of the two pipes.
//   cnt trick: mk is 0xFFFF per active half.  word -= mk adds 1 to the low count and (hi_active - lo_active) to the high one;
//   the caller repairs the high half once per row with word += word << 16.
__device__ __forceinline__ void add_relu_pair(uint32_t q2, uint32_t p2, float& a_lo, float& a_hi, uint32_t& cn, uint32_t& mout,
                                              uint32_t bitc) {
  uint32_t rb;
  asm("fma.rn.relu.bf16x2 %0, %1, %2, %3;" : "=r"(rb) : "r"(p2), "r"(0x3f803f80u), "r"(q2));
  const __nv_bfloat162 zero = __floats2bfloat162_rn(0.f, 0.f);
  const uint32_t mk = __hgt2_mask(*reinterpret_cast<const __nv_bfloat162*>(&rb), zero);  // relu(s) > 0  <=>  s > 0
  fadd_bf16x2(rb, a_lo, a_hi);
  cn -= mk;
  mout |= mk & bitc;
}

// acc[2c], acc[2c + 1] += the halves of word c that the mask byte selects: predicated FHADD.BF16, predicates from one R2P
__device__ __forceinline__ void masked_add8(const uint4& v, uint32_t mbyte, float* acc) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    unsigned short l, h;
    asm("mov.b32 {%0, %1}, %2;" : "=h"(l), "=h"(h) : "r"(w[c]));
    if (mbyte & (1u << (2 * c))) asm("add.rn.f32.bf16 %0, %1, %0;" : "+f"(acc[2 * c]) : "h"(l));
    if (mbyte & (2u << (2 * c))) asm("add.rn.f32.bf16 %0, %1, %0;" : "+f"(acc[2 * c + 1]) : "h"(h));
  }
}

// One edge worth of work for one lane chunk (8 columns).
template <int MODE>
__device__ __forceinline__ void seg_accumulate(const uint4& v, const uint4& v2, const float* rop, const uint4& rop2, float w,
                                               bool weighted, uint32_t mbyte, float* acc, uint32_t* cn2, uint32_t& mout) {
  if (MODE == SEG_ADD_RELU) {
    uint32_t m = 0;
    add_relu_pair(v.x, rop2.x, acc[0], acc[1], cn2[0], m, 0x00020001u);
    add_relu_pair(v.y, rop2.y, acc[2], acc[3], cn2[1], m, 0x00080004u);
    add_relu_pair(v.z, rop2.z, acc[4], acc[5], cn2[2], m, 0x00200010u);
    add_relu_pair(v.w, rop2.w, acc[6], acc[7], cn2[3], m, 0x00800040u);
    mout = m | (m >> 16);  // only the low byte is stored
    return;
  }
  if (MODE == SEG_MASKED_SUM) {
    masked_add8(v, mbyte, acc);
    return;
  }
  if (MODE == SEG_SUM && !weighted) {  // plain / mean aggregation: widening adds straight from the packed words
    fadd_bf16x8(v, acc);
    return;
  }
  float f[8];
  unpack8(v, f);
  if (MODE == SEG_SUM) {
#pragma unroll
    for (int t = 0; t < 8; ++t) acc[t] = fmaf(w, f[t], acc[t]);
  } else {  // SEG_ADD_RELU_BWD: f = dOut[i], f2 = P[i], rop = Q[j]
    float f2[8];
    unpack8(v2, f2);
#pragma unroll
    for (int t = 0; t < 8; ++t)
      if (f2[t] + rop[t] > 0.f) acc[t] += f[t];
  }
}

// One output row (segment) handled by a group of G lanes.  FULL: the row has exactly G*NC 8-column chunks, so every lane
// owns valid columns and the chunk count is a compile-time constant (no per-edge column predicates, immediate offsets for
// the mask bytes).  AUX: how the ReLU bit mask is addressed -- 0 = not used, 1 = rows in this walk's slot order (running
// pointer, streaming), 2 = rows through p.slot_map (one extra index load + shuffle per edge).  SEG_ADD_RELU writes the mask,
// SEG_MASKED_SUM reads it; exactly one of the two walks can be sequential, the other one pays the random 32-byte accesses
// (writes are fire-and-forget, reads stall: see DESIGN.md).  FULL and AUX are block-uniform run-time facts that the
// kernel turns into template arguments once, outside the edge loop (ncu/SASS: the per-edge null check of p.mask cost two
// uniform compares, a constant reload and two R2UR per edge because it evicted the L2-policy descriptor).
//
// The lane group walks rows r, r + stride, ... < r_end and software-pipelines the dependent index loads -- while row r is being
// reduced, the segment bounds of row r + 2*stride and the first G neighbour indices of row r + stride are already in flight.
// Without this every row (25 edges on average for the products graph) pays three serialised memory latencies
// (ptr -> idx -> gather) before its first useful instruction.  The kernel gives each group a short run of CONSECUTIVE rows
// (p.rows_per_group) and keeps the grid non-persistent: a grid-strided persistent variant was measured 35 % slower because the
// hardware block scheduler otherwise keeps all SMs inside one compact window of rows, which is what makes the evict_last L2
// policy on the gathered rows pay.
template <int MODE, int G, int NC, bool HALO, bool FULL, int AUX, bool TAILPF>
__device__ __forceinline__ void seg_rows(const SegParams& p, const char* const* s_peer, int r, const int stride, const int n_rows,
                                         const int g, const unsigned gmask) {
  constexpr int UNROLL = 4;
  constexpr bool PF = TAILPF && HALO && NC == 1 && (MODE == SEG_ADD_RELU || MODE == SEG_SUM);  // prefetch the tail (remote) edges of a row
  constexpr bool USES_MASK = (MODE == SEG_ADD_RELU && AUX != 0) || MODE == SEG_MASKED_SUM;
  constexpr bool MAPPED = USES_MASK && AUX == 2;
  const int nchunk = FULL ? (G * NC) : (p.H >> 3);  // 8-column chunks per row
  const uint64_t pol_keep = l2_policy_evict_last();
  const uint64_t pol_stream = l2_policy_evict_first();
  const bool weighted = (MODE == SEG_SUM) && p.slot_w != nullptr;

  // pipeline registers: current row (beg, end, first-group indices) and the bounds of the next one
  int beg = p.ptr[r], end = p.ptr_end[r];
  int cur_j = 0, cur_s = 0;
  float cur_w = 1.f;
  if (g < min(G, end - beg)) {
    cur_j = p.idx[beg + g];
    if (MODE == SEG_SUM && p.slot_w) cur_w = p.slot_w[beg + g];
    if (MAPPED) cur_s = p.slot_map[beg + g];
  }
  int rn = r + stride;
  int nbeg = 0, nend = 0;
  if (rn < n_rows) { nbeg = p.ptr[rn]; nend = p.ptr_end[rn]; }

  while (true) {
  // ---- issue the prefetches for the following rows (consumed after this row's reduction) ----
  int nxt_j = 0, nxt_s = 0;
  float nxt_w = 1.f;
  if (g < min(G, nend - nbeg)) {
    nxt_j = p.idx[nbeg + g];
    if (MODE == SEG_SUM && p.slot_w) nxt_w = p.slot_w[nbeg + g];
    if (MAPPED) nxt_s = p.slot_map[nbeg + g];
  }
  const int rnn = rn + stride;
  int nnbeg = 0, nnend = 0;
  if (rnn < n_rows) { nnbeg = p.ptr[rnn]; nnend = p.ptr_end[rnn]; }

  const bool accum = (MODE == SEG_ADD_RELU) && p.accumulate != 0;
  const bool skip_out2 = (MODE == SEG_MASKED_SUM || MODE == SEG_SUM) && p.out2_skip_empty && beg == end && r >= p.out_split;
  if (!(accum && beg == end) && !skip_out2) {  // an accumulate pass leaves rows without edges in its range untouched
  float acc[NC][8];
  float rop[NC][8];
  uint4 rop2[NC];  // SEG_ADD_RELU: the row operand stays packed (bf16x2)
  uint32_t cn2[NC][4];
  bool colok[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int ch = g + c * G;
    colok[c] = FULL ? true : (ch < nchunk);
#pragma unroll
    for (int t = 0; t < 8; ++t) { acc[c][t] = 0.f; rop[c][t] = 0.f; }
#pragma unroll
    for (int t = 0; t < 4; ++t) cn2[c][t] = 0;
    rop2[c] = make_uint4(0, 0, 0, 0);
    if (accum && colok[c]) {  // continue from the first pass: out row back to fp32, counters back to the borrow-free form
      unpack8(ldg_v4(p.out + (int64_t)r * p.ld_out + ch * 8), acc[c]);
      if (p.cnt) {
        const uint4 cv = ldg_v4(p.cnt + (int64_t)r * p.ld_cnt + ch * 8);
        cn2[c][0] = cv.x - (cv.x << 16); cn2[c][1] = cv.y - (cv.y << 16);
        cn2[c][2] = cv.z - (cv.z << 16); cn2[c][3] = cv.w - (cv.w << 16);
      }
    }
    if (MODE == SEG_ADD_RELU && colok[c]) rop2[c] = ldg_hint_v4(p.rowop + (int64_t)r * p.ld_rowop + ch * 8, pol_stream);
    if (MODE == SEG_ADD_RELU_BWD && colok[c])
      unpack8(ldg_hint_v4(p.rowop + (int64_t)r * p.ld_rowop + ch * 8, pol_stream), rop[c]);
  }
  const char* src_lane = reinterpret_cast<const char*>(p.src) + g * 16;
  const char* src2_lane = reinterpret_cast<const char*>(p.src2) + g * 16;

  // SEG_ADD_RELU + AUX: running pointer to this lane's mask byte of the current edge slot
  unsigned char* mp = (USES_MASK && !MAPPED) ? p.mask + (int64_t)beg * nchunk + g : nullptr;  // running pointer
  unsigned char* mask_lane = MAPPED ? p.mask + g : nullptr;
  for (int base = beg; base < end; base += G) {
    const int n = min(G, end - base);
    int my_j = cur_j, my_s = cur_s;  // first group: prefetched while the previous row was reduced
    float my_w = cur_w;
    if (base != beg) {
      my_j = 0; my_s = 0; my_w = 1.f;
      if (g < n) {
        my_j = p.idx[base + g];
        if (MODE == SEG_SUM && p.slot_w) my_w = p.slot_w[base + g];
        if (MAPPED) my_s = p.slot_map[base + g];
      }
    }
    // Pull kernels: the remote neighbours of a row sort behind its local ones, i.e. into the tail of its last chunk.  Their loads are
    // issued here, before the local neighbours are processed, and consumed after them -- in the same order as without the prefetch,
    // so results are bit-identical -- which gives the NVLink round trip (3-4x a local L2 hit) the whole row to complete.
    int n_main = n, nt = 0;
    uint4 vt[PF ? 4 : 1][NC];
    float wt[PF ? 4 : 1];
    int st[PF ? 4 : 1];
    if (PF && base + G >= end) {
      nt = min(4, n);
      n_main = n - nt;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (u < nt) {
          const int j = __shfl_sync(gmask, my_j, n_main + u, G);
          wt[u] = (MODE == SEG_SUM) ? __shfl_sync(gmask, my_w, n_main + u, G) : 1.f;
          st[u] = MAPPED ? __shfl_sync(gmask, my_s, n_main + u, G) : 0;
          const char* row = src_row<HALO>(p, s_peer, src_lane, g * 16, j);
#pragma unroll
          for (int c = 0; c < NC; ++c)
            if (colok[c]) vt[u][c] = ldg_any_hint_v4(row + c * G * 16, pol_keep);
        }
      }
    }
    int t0 = 0;
    for (; t0 + UNROLL <= n_main; t0 += UNROLL) {  // full groups: no per-edge predication
      uint4 v[UNROLL][NC], v2[UNROLL][NC];
      uint32_t mb[UNROLL][NC];
      float w[UNROLL];
      int sl[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const int j = __shfl_sync(gmask, my_j, t0 + u, G);
        if (MODE == SEG_SUM) w[u] = __shfl_sync(gmask, my_w, t0 + u, G); else w[u] = 1.f;
        sl[u] = MAPPED ? __shfl_sync(gmask, my_s, t0 + u, G) : 0;
        const char* row = src_row<HALO>(p, s_peer, src_lane, g * 16, j);
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          if (colok[c]) {
            v[u][c] = HALO ? ldg_any_hint_v4(row + c * G * 16, pol_keep) : ldg_hint_v4(row + c * G * 16, pol_keep);
            if (MODE == SEG_ADD_RELU_BWD) v2[u][c] = ldg_nc_v4(src2_lane + (uint64_t)(uint32_t)j * p.ld_src2_b + c * G * 16);
            if (MODE == SEG_MASKED_SUM)
              mb[u][c] = MAPPED ? __ldg(mask_lane + (int64_t)sl[u] * nchunk + c * G) : ldg_stream_u8(mp + u * nchunk + c * G);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          if (colok[c]) {
            uint32_t mout = 0;
            seg_accumulate<MODE>(v[u][c], v2[u][c], rop[c], rop2[c], w[u], weighted, mb[u][c], acc[c], cn2[c], mout);
            if (MODE == SEG_ADD_RELU && AUX == 1) stg_cs_u8(mp + u * nchunk + c * G, mout);
            if (MODE == SEG_ADD_RELU && AUX == 2) stg_cs_u8(mask_lane + (int64_t)sl[u] * nchunk + c * G, mout);
          }
        }
      }
      if (USES_MASK && !MAPPED) mp += UNROLL * nchunk;
    }
    for (; t0 < n_main; ++t0) {  // tail
      const int j = __shfl_sync(gmask, my_j, t0, G);
      const float w = (MODE == SEG_SUM) ? __shfl_sync(gmask, my_w, t0, G) : 1.f;
      const int sl = MAPPED ? __shfl_sync(gmask, my_s, t0, G) : 0;
      const char* row = src_row<HALO>(p, s_peer, src_lane, g * 16, j);
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        if (colok[c]) {
          const uint4 v = HALO ? ldg_any_hint_v4(row + c * G * 16, pol_keep) : ldg_hint_v4(row + c * G * 16, pol_keep);
          uint4 v2 = make_uint4(0, 0, 0, 0);
          uint32_t mb = 0, mout = 0;
          if (MODE == SEG_ADD_RELU_BWD) v2 = ldg_nc_v4(src2_lane + (uint64_t)(uint32_t)j * p.ld_src2_b + c * G * 16);
          if (MODE == SEG_MASKED_SUM) mb = MAPPED ? __ldg(mask_lane + (int64_t)sl * nchunk + c * G) : ldg_stream_u8(mp + c * G);
          seg_accumulate<MODE>(v, v2, rop[c], rop2[c], w, weighted, mb, acc[c], cn2[c], mout);
          if (MODE == SEG_ADD_RELU && AUX == 1) stg_cs_u8(mp + c * G, mout);
          if (MODE == SEG_ADD_RELU && AUX == 2) stg_cs_u8(mask_lane + (int64_t)sl * nchunk + c * G, mout);
        }
      }
      if (USES_MASK && !MAPPED) mp += nchunk;
    }
    if (PF) {  // the prefetched tail edges, in slot order
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (u < nt) {
#pragma unroll
          for (int c = 0; c < NC; ++c) {
            if (colok[c]) {
              uint32_t mout = 0;
              seg_accumulate<MODE>(vt[u][c], vt[u][c], rop[c], rop2[c], wt[u], weighted, 0u, acc[c], cn2[c], mout);
              if (MODE == SEG_ADD_RELU && AUX == 1) stg_cs_u8(mp + c * G, mout);
              if (MODE == SEG_ADD_RELU && AUX == 2) stg_cs_u8(mask_lane + (int64_t)st[u] * nchunk + c * G, mout);
            }
          }
          if (USES_MASK && !MAPPED) mp += nchunk;
        }
      }
    }
  }

  const float scale = (MODE == SEG_SUM && p.mean) ? 1.f / (float)max(end - beg, 1) : 1.f;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    if (colok[c]) {
      const int colo = (g + c * G) * 8;
      if (MODE == SEG_SUM && p.mean) {
#pragma unroll
        for (int t = 0; t < 8; ++t) acc[c][t] *= scale;
      }
      __nv_bfloat16* orow = p.out + (int64_t)r * p.ld_out;
      if ((MODE == SEG_MASKED_SUM || MODE == SEG_SUM) && r >= p.out_split) orow = p.out2 + (int64_t)(r - p.out_split) * p.ld_out2;
      stg_cs_v4(orow + colo, pack8(acc[c]));
      if (MODE == SEG_MASKED_SUM && p.dp_out && r < p.dp_rows) {
        float gr[8];
        unpack8(ldg_nc_v4(reinterpret_cast<const char*>(p.src) + (uint64_t)r * p.ld_src_b + colo * 2), gr);
        const uint4 cv = ldg_nc_v4(p.cnt + (int64_t)r * p.ld_cnt + colo);
        gr[0] *= (float)(cv.x & 0xffff); gr[1] *= (float)(cv.x >> 16);
        gr[2] *= (float)(cv.y & 0xffff); gr[3] *= (float)(cv.y >> 16);
        gr[4] *= (float)(cv.z & 0xffff); gr[5] *= (float)(cv.z >> 16);
        gr[6] *= (float)(cv.w & 0xffff); gr[7] *= (float)(cv.w >> 16);
        stg_cs_v4(p.dp_out + (int64_t)r * p.ld_dp + colo, pack8(gr));
      }
      if (MODE == SEG_ADD_RELU && p.cnt) {
#pragma unroll
        for (int t = 0; t < 4; ++t) cn2[c][t] += cn2[c][t] << 16;  // repair the high halves (see add_relu_pair)
        uint4 cv = make_uint4(cn2[c][0], cn2[c][1], cn2[c][2], cn2[c][3]);
        stg_cs_v4(p.cnt + (int64_t)r * p.ld_cnt + colo, cv);
      }
    }
  }

  }  // !(accum && empty range)

  // ---- rotate the pipeline ----
  if (rn >= n_rows) break;
  r = rn; beg = nbeg; end = nend;
  cur_j = nxt_j; cur_s = nxt_s; cur_w = nxt_w;
  rn = rnn; nbeg = nnbeg; nend = nnend;
  }  // while rows
}

template <int MODE, int G, int NC, bool HALO, bool TAILPF>
__global__ void __launch_bounds__(256, (NC == 1) ? (TAILPF ? 3 : (MODE == SEG_SUM ? 5 : 4)) : 1) seg_reduce_kernel(const SegParams p) {
  __shared__ const char* s_peer[HALO ? kMaxPeers : 1];
  if (HALO) {
    if ((int)threadIdx.x < p.world) s_peer[threadIdx.x] = reinterpret_cast<const char*>(p.peer_base[threadIdx.x]);
    __syncthreads();
  }
  const int lane = threadIdx.x & 31;
  const int g = lane & (G - 1);
  const int sub = lane / G;
  const int warp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int r = (warp * (32 / G) + sub) * p.rows_per_group;  // first of this group's consecutive rows
  const int stride = 1;
  if (r >= p.n_rows) return;
  const int r_end = min(r + p.rows_per_group, (int)p.n_rows);
  const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << (sub * G));
  const bool full = (p.H >> 3) == G * NC;
  if (MODE == SEG_ADD_RELU || MODE == SEG_MASKED_SUM) {
    const int aux = p.mask == nullptr ? 0 : (p.slot_map ? 2 : 1);
    if (full) {
      if (aux == 2) seg_rows<MODE, G, NC, HALO, true, 2, TAILPF>(p, s_peer, r, stride, r_end, g, gmask);
      else if (aux == 1) seg_rows<MODE, G, NC, HALO, true, 1, TAILPF>(p, s_peer, r, stride, r_end, g, gmask);
      else if (MODE == SEG_ADD_RELU) seg_rows<MODE, G, NC, HALO, true, 0, TAILPF>(p, s_peer, r, stride, r_end, g, gmask);
    } else {
      if (aux == 2) seg_rows<MODE, G, NC, HALO, false, 2, TAILPF>(p, s_peer, r, stride, r_end, g, gmask);
      else if (aux == 1) seg_rows<MODE, G, NC, HALO, false, 1, TAILPF>(p, s_peer, r, stride, r_end, g, gmask);
      else if (MODE == SEG_ADD_RELU) seg_rows<MODE, G, NC, HALO, false, 0, TAILPF>(p, s_peer, r, stride, r_end, g, gmask);
    }
  } else {
    if (full) seg_rows<MODE, G, NC, HALO, true, 0, TAILPF>(p, s_peer, r, stride, r_end, g, gmask);
    else seg_rows<MODE, G, NC, HALO, false, 0, TAILPF>(p, s_peer, r, stride, r_end, g, gmask);
  }
}

template <int MODE, bool HALO>
static int launch_seg(const SegParams& p, cudaStream_t s) {
  if (p.n_rows == 0) return 0;
  const int chunks = p.H / 8;
  const int warps_per_block = 8;
  if (p.n_rows >= (int64_t)1 << 31) return (int)cudaErrorInvalidValue;
  static int rows_per_group = 0;
  if (rows_per_group == 0) {
    const char* e = getenv("DGRAPH_B200_SEG_ROWS");
    rows_per_group = e ? atoi(e) : 8;
    if (rows_per_group < 1 || rows_per_group > 64) rows_per_group = 8;
  }
  // DGRAPH_B200_PULL_PREFETCH=1: pull kernels issue the loads of a row's tail (remote) edges before its local ones.  Measured on
  // 8 x B200 (products-shaped GCN step): 3.94 ms with, 3.61 ms without -- the 16 extra live registers cost a resident block per SM
  // (80 vs 64 registers), which hurts the local gathers more than the earlier NVLink loads help.  Off by default.
  const char* pfe = HALO ? getenv("DGRAPH_B200_PULL_PREFETCH") : nullptr;
  const bool tail_prefetch = HALO && pfe && pfe[0] == '1';
  SegParams q = p;
  q.rows_per_group = rows_per_group;
  if (q.ptr_end == nullptr) q.ptr_end = q.ptr + 1;
  if (q.out2 == nullptr) q.out_split = q.n_rows;
#define DGB_LAUNCH(G, NC)                                                                 \
  do {                                                                                    \
    const int64_t rows_per_block = (int64_t)warps_per_block * (32 / (G)) * rows_per_group; \
    const int64_t blocks = (p.n_rows + rows_per_block - 1) / rows_per_block;              \
    if (HALO && (NC) == 1 && tail_prefetch)                                               \
      seg_reduce_kernel<MODE, G, NC, HALO, HALO && (NC) == 1><<<(unsigned)blocks, warps_per_block * 32, 0, s>>>(q); \
    else                                                                                  \
      seg_reduce_kernel<MODE, G, NC, HALO, false><<<(unsigned)blocks, warps_per_block * 32, 0, s>>>(q); \
  } while (0)
  if (chunks <= 4) DGB_LAUNCH(4, 1);
  else if (chunks <= 8) DGB_LAUNCH(8, 1);
  else if (chunks <= 16) DGB_LAUNCH(16, 1);
  else if (chunks <= 32) DGB_LAUNCH(32, 1);
  else if (chunks <= 64) DGB_LAUNCH(32, 2);
  else if (chunks <= 128) DGB_LAUNCH(32, 4);
  else return (int)cudaErrorInvalidValue;
#undef DGB_LAUNCH
  DGB_CHECK_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------------------------------------------
// elementwise: dP = dOut * cnt  (backward of the row-operand of SEG_ADD_RELU)
// ---------------------------------------------------------------------------------------------------
__global__ void mul_cnt_kernel(const __nv_bfloat16* __restrict__ g, int64_t ld_g, const unsigned short* __restrict__ cnt,
                               int64_t ld_c, __nv_bfloat16* __restrict__ out, int64_t ld_o, int64_t rows, int H8) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * H8) return;
  const int64_t r = i / H8;
  const int c = (int)(i % H8) * 8;
  uint4 gv = ldg_nc_v4(g + r * ld_g + c);
  uint4 cv = ldg_nc_v4(cnt + r * ld_c + c);
  float f[8];
  unpack8(gv, f);
  f[0] *= (float)(cv.x & 0xffff); f[1] *= (float)(cv.x >> 16);
  f[2] *= (float)(cv.y & 0xffff); f[3] *= (float)(cv.y >> 16);
  f[4] *= (float)(cv.z & 0xffff); f[5] *= (float)(cv.z >> 16);
  f[6] *= (float)(cv.w & 0xffff); f[7] *= (float)(cv.w >> 16);
  stg_v4(out + r * ld_o + c, pack8(f));
}

}  // namespace dgb

using namespace dgb;

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static inline bool fits32(int64_t v) { return v >= 0 && v < (int64_t(1) << 32); }

// out[r,:] = (mean ? 1/deg : 1) * sum_{k in seg r} w[k] * src[idx[k],:]          (bf16 in/out, fp32 accumulate)
// Pull variant (peer_base != null): idx[k] < 0 names a remote row, ~idx[k] = (owner << peer_shift) | row, fetched from
// peer_base[owner] + row * ld_peer inside the kernel.  out2 (optional): rows >= out_split are written to out2[r - out_split].
static inline bool peer_args_ok(const void* const* peer_base, int world, int peer_shift, int64_t ld_peer) {
  if (!peer_base) return true;
  return world >= 1 && world <= kMaxPeers && peer_shift >= 1 && peer_shift <= 30 && ((int64_t)world << peer_shift) <= ((int64_t)1 << 31) &&
         ld_peer % 8 == 0 && fits32(ld_peer * 2);
}

DGB_API int dgb_seg_sum(const int* ptr, const int* idx, int64_t n_rows, int H, const void* src, int64_t ld_src,
                        const float* slot_w, int mean, void* out, int64_t ld_out, const void* const* peer_base, int world,
                        int peer_shift, int64_t ld_peer, void* out2, int64_t ld_out2, int64_t out_split, void* stream) {
  if (H % 8 || ld_src % 8 || ld_out % 8 || !aligned16(src) || !aligned16(out) || (out2 && (ld_out2 % 8 || !aligned16(out2))))
    return (int)cudaErrorInvalidValue;
  SegParams p{};
  p.ptr = ptr; p.idx = idx; p.n_rows = n_rows; p.H = H;
  if (!fits32(ld_src * 2) || !peer_args_ok(peer_base, world, peer_shift, ld_peer)) return (int)cudaErrorInvalidValue;
  p.src = (const __nv_bfloat16*)src; p.ld_src_b = (uint32_t)(ld_src * 2);
  p.slot_w = slot_w; p.mean = mean;
  p.out = (__nv_bfloat16*)out; p.ld_out = ld_out;
  p.out2 = (__nv_bfloat16*)out2; p.ld_out2 = ld_out2; p.out_split = out_split;
  p.peer_base = (const __nv_bfloat16* const*)peer_base; p.world = world; p.peer_shift = peer_shift;
  p.ld_peer_b = (uint32_t)(ld_peer * 2);
  return peer_base ? launch_seg<SEG_SUM, true>(p, (cudaStream_t)stream) : launch_seg<SEG_SUM, false>(p, (cudaStream_t)stream);
}

// out[r,:] = sum_{k in seg r} relu(P[r,:] + Q[idx[k],:]);   cnt[r,c] = #active edges (optional)
// mask_slot_map (optional): mask row of edge slot k is mask_slot_map[k] instead of k -- pass the CSR->CSC slot permutation so
// that the backward CSC walk (dgb_seg_masked_sum with slot_map = null) streams the mask rows sequentially.
// ptr_end (optional, null = ptr + 1) and accumulate: walk only slots [ptr[r], ptr_end[r]) of every row and, with accumulate = 1,
// add to the out / cnt rows an earlier launch produced.
// Pull variant (peer_base != null): Q rows with idx[k] < 0 are read from the owner GPU inside the kernel (see dgb_seg_sum); this is
// the fused halo exchange + aggregation of a distributed GraphConv layer (no halo buffer, no exchange kernel).
DGB_API int dgb_seg_add_relu_fwd_ex(const int* ptr, const int* ptr_end, const int* idx, int64_t n_rows, int H, const void* P,
                                    int64_t ld_p, const void* Q, int64_t ld_q, void* out, int64_t ld_out, unsigned short* cnt,
                                    int64_t ld_cnt, unsigned char* mask, const int* mask_slot_map, int accumulate,
                                    const void* const* peer_base, int world, int peer_shift, int64_t ld_peer, void* stream) {
  if (H % 8 || ld_p % 8 || ld_q % 8 || ld_out % 8 || (cnt && ld_cnt % 8) || !aligned16(P) || !aligned16(Q) ||
      !aligned16(out))
    return (int)cudaErrorInvalidValue;
  SegParams p{};
  p.ptr = ptr; p.ptr_end = ptr_end; p.accumulate = accumulate; p.idx = idx; p.n_rows = n_rows; p.H = H;
  if (!fits32(ld_q * 2) || !peer_args_ok(peer_base, world, peer_shift, ld_peer)) return (int)cudaErrorInvalidValue;
  p.src = (const __nv_bfloat16*)Q; p.ld_src_b = (uint32_t)(ld_q * 2);
  p.rowop = (const __nv_bfloat16*)P; p.ld_rowop = ld_p;
  p.out = (__nv_bfloat16*)out; p.ld_out = ld_out; p.cnt = cnt; p.ld_cnt = ld_cnt; p.mask = mask;
  p.slot_map = mask ? mask_slot_map : nullptr;
  p.peer_base = (const __nv_bfloat16* const*)peer_base; p.world = world; p.peer_shift = peer_shift;
  p.ld_peer_b = (uint32_t)(ld_peer * 2);
  return peer_base ? launch_seg<SEG_ADD_RELU, true>(p, (cudaStream_t)stream)
                   : launch_seg<SEG_ADD_RELU, false>(p, (cudaStream_t)stream);
}

DGB_API int dgb_seg_add_relu_fwd(const int* ptr, const int* idx, int64_t n_rows, int H, const void* P, int64_t ld_p,
                                 const void* Q, int64_t ld_q, void* out, int64_t ld_out, unsigned short* cnt,
                                 int64_t ld_cnt, unsigned char* mask, void* stream) {
  return dgb_seg_add_relu_fwd_ex(ptr, nullptr, idx, n_rows, H, P, ld_p, Q, ld_q, out, ld_out, cnt, ld_cnt, mask, nullptr, 0,
                                 nullptr, 0, 0, 0, stream);
}

// dQ[j,:] = sum_{k in CSC seg j} (P[row[k],:] + Q[j,:] > 0) ? dOut[row[k],:] : 0       (transposed walk)
DGB_API int dgb_seg_add_relu_bwd(const int* colptr, const int* row, int64_t n_aug, int H, const void* dOut,
                                 int64_t ld_g, const void* P, int64_t ld_p, const void* Q, int64_t ld_q, void* dQ,
                                 int64_t ld_dq, void* stream) {
  if (H % 8 || ld_g % 8 || ld_p % 8 || ld_q % 8 || ld_dq % 8) return (int)cudaErrorInvalidValue;
  SegParams p{};
  p.ptr = colptr; p.idx = row; p.n_rows = n_aug; p.H = H;
  if (!fits32(ld_g * 2) || !fits32(ld_p * 2)) return (int)cudaErrorInvalidValue;
  p.src = (const __nv_bfloat16*)dOut; p.ld_src_b = (uint32_t)(ld_g * 2);
  p.src2 = (const __nv_bfloat16*)P; p.ld_src2_b = (uint32_t)(ld_p * 2);
  p.rowop = (const __nv_bfloat16*)Q; p.ld_rowop = ld_q;
  p.out = (__nv_bfloat16*)dQ; p.ld_out = ld_dq;
  return launch_seg<SEG_ADD_RELU_BWD, false>(p, (cudaStream_t)stream);
}

// dQ[j,:] = sum_{k in CSC seg j} mask[slot_map[k]] ? dOut[row[k],:] : 0, with the ReLU bit masks the forward kernel saved
// (one byte per edge slot and 8-column chunk): one row gather per edge instead of two, no recomputation.  slot_map = null
// means the mask rows are already in this walk's slot order (forward used dgb_seg_add_relu_fwd_mapped).  When dP is given,
// rows j < dp_rows also write dP[j,:] = dOut[j,:] * cnt[j,:] in the same pass.  colptr_end (optional, null = colptr + 1): segment j is
// [colptr[j], colptr_end[j]) -- lets a launch walk the chunks of split hub columns as rows of their own.  dQ2 (optional): rows j >= dq_split (the halo rows of
// a distributed layer) are written to dQ2[j - dq_split, :] -- the symmetric slot their owners pull-reduce from; with dq2_skip_empty
// those of them that have no edge in this (restricted) walk are not written at all (the owners' restricted plan never reads them).
DGB_API int dgb_seg_masked_sum(const int* colptr, const int* colptr_end, const int* row, const int* slot_map,
                               const unsigned char* mask, int64_t n_aug, int H, const void* dOut, int64_t ld_g, void* dQ, int64_t ld_dq,
                               const unsigned short* cnt, int64_t ld_cnt, void* dP, int64_t ld_dp, int64_t dp_rows,
                               void* dQ2, int64_t ld_dq2, int64_t dq_split, int dq2_skip_empty, void* stream) {
  if (H % 8 || ld_g % 8 || ld_dq % 8 || !fits32(ld_g * 2) || !mask || (dP && (ld_dp % 8 || ld_cnt % 8 || !cnt)) ||
      (dQ2 && (ld_dq2 % 8 || !aligned16(dQ2))))
    return (int)cudaErrorInvalidValue;
  SegParams p{};
  p.ptr = colptr; p.ptr_end = colptr_end; p.idx = row; p.n_rows = n_aug; p.H = H;
  p.src = (const __nv_bfloat16*)dOut; p.ld_src_b = (uint32_t)(ld_g * 2);
  p.mask = const_cast<unsigned char*>(mask); p.slot_map = slot_map;
  p.out = (__nv_bfloat16*)dQ; p.ld_out = ld_dq;
  p.cnt = const_cast<unsigned short*>(cnt); p.ld_cnt = ld_cnt;
  p.dp_out = (__nv_bfloat16*)dP; p.ld_dp = ld_dp; p.dp_rows = dP ? dp_rows : 0;
  p.out2 = (__nv_bfloat16*)dQ2; p.ld_out2 = ld_dq2; p.out_split = dq_split; p.out2_skip_empty = dQ2 ? dq2_skip_empty : 0;
  return launch_seg<SEG_MASKED_SUM, false>(p, (cudaStream_t)stream);
}

// dP = dOut * cnt
DGB_API int dgb_mul_cnt(const void* g, int64_t ld_g, const unsigned short* cnt, int64_t ld_c, void* out, int64_t ld_o,
                        int64_t rows, int H, void* stream) {
  if (H % 8 || ld_g % 8 || ld_c % 8 || ld_o % 8) return (int)cudaErrorInvalidValue;
  const int64_t n = rows * (H / 8);
  if (n == 0) return 0;
  mul_cnt_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)g, ld_g, cnt, ld_c, (__nv_bfloat16*)out, ld_o, rows, H / 8);
  DGB_CHECK_LAUNCH();
  return 0;
}
