// Batched Hamming / euclid_lsh scan with exact top-k  (sm_100a).
//
// Replaces, for a batch of queries,
//   [core] nearest_neighbor/bit_vector_ranking.cpp  ranking_hamming_bit_vectors
//   [core] nearest_neighbor/euclid_lsh.cpp          neighbor_row_from_hash
//   [core] storage/fixed_size_heap.hpp
// (jubatus_core 0.3.2, reached from jubatus/server/server/nearest_neighbor_serv.cpp:144,153,
//  162,171).  Result order is the reference's: ascending (distance, row index).
//
// One CTA streams a contiguous range of 32-row tiles once and ranks `qp` queries against
// it.  Lane = row in tile; the row's words sit in registers, query words are broadcast
// from shared memory, XOR + popcount (optionally through a carry-save adder tree that
// trades POPC-pipe work for LOP3s).  Each (CTA, query) owns a candidate list in shared
// memory guarded by a threshold: a row is appended only if its key
// ((distance << 32) | row) is below the k-th smallest key seen so far by the CTA, so after
// a short warm-up almost every row costs one compare.  A full list is compacted (bitonic
// sort by the appending warp, keep k, tighten the threshold) under a per-query spin lock.
// The CTA's k best are written to scratch; merge_local_kernel reduces the per-CTA lists.
#include "internal.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

namespace jbn {

#define FULL_MASK 0xffffffffu
static const int SCAN_THREADS = 512;
static const int SCAN_WARPS = SCAN_THREADS / 32;
static const int MERGE_THREADS = 256;
static const uint32_t MAX_CAP = 4096;  // => at most 2048 results per pass

struct ScanArgs {
  const unsigned char* slab;
  const float* norms;
  uint64_t local_rows;
  uint64_t ntiles;
  const uint64_t* qsigs;
  const float* qnorms;
  const double* coslut;
  const uint64_t* floor_keys;
  uint32_t* ghist;            // [nq] shared bound words, then [nq][hist_stride] bin counts (nullable)
  uint32_t hist_stride;
  const uint64_t* tau_init;   // per query: exclusive upper bound of admissible keys from the sample pre-pass (nullable)
  uint64_t* cand;
  uint32_t nq, k, cap, qp, w64, hash_num, tiles_per_chunk, nchunks;
  uint32_t w32;
  uint32_t interleave;  // 1: CTAs take turns tile-group by tile-group (one sweep front); 0: one contiguous range per CTA
  unsigned long long* trace;   // tuning experiments ($JUBANN_SCAN_TRACE): per (CTA < 4, group) timestamps of warp 0
  uint32_t score_mode;  // euclid_lsh: 0 = score in double (cos table of doubles), 1 = all in float (table holds cosf values)
  uint32_t debug;       // tuning experiments ($JUBANN_SCAN_DEBUG): 1 no early refresh, 2 no shared histogram, 4 no polling, 8 no refresh after a flush
};

// ---------------------------------------------------------------------------
// streaming loads: read-once data, keep it out of L1
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint4 ld_stream_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint2 ld_stream_v2(const void* p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ float ld_stream_f32(const float* p) {
  float r;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}

// the same addresses as a hint: bring the tile into L2 while the current one is being ranked
template <int W32>
__device__ __forceinline__ void prefetch_row_l2(const unsigned char* slab, uint64_t tile, int lane) {
  if (W32 == 2) {
    asm volatile("prefetch.global.L2 [%0];" ::"l"(slab + (tile * 32 + lane) * 8ull));
  } else {
    constexpr int C = W32 / 4;
#pragma unroll
    for (int c = 0; c < C; ++c) asm volatile("prefetch.global.L2 [%0];" ::"l"(slab + ((tile * C + c) * 32 + lane) * 16ull));
  }
}

template <int W32>
__device__ __forceinline__ void load_row(const unsigned char* slab, uint64_t tile, int lane,
                                         uint32_t (&r)[W32]) {
  if (W32 == 2) {
    uint2 v = ld_stream_v2(slab + (tile * 32 + lane) * 8ull);
    r[0] = v.x; r[1] = v.y;
  } else {
    constexpr int C = W32 / 4;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      uint4 v = ld_stream_v4(slab + ((tile * C + c) * 32 + lane) * 16ull);
      r[4 * c + 0] = v.x; r[4 * c + 1] = v.y; r[4 * c + 2] = v.z; r[4 * c + 3] = v.w;
    }
  }
}

// ---------------------------------------------------------------------------
// Hamming distance of the register row against a query in shared memory.
// PV 0: one POPC per word.  PV 1/2: carry-save adders (3 words -> sum + carry, two LOP3)
// so fewer words reach the quarter-rate POPC pipe.
// ---------------------------------------------------------------------------
// Written as the two LOP3 themselves (sum = 0x96, carry = majority 0xE8).  Left as C expressions,
// the compiler re-derives sum and carry from the six raw operands (row and query words) when
// several rows and queries share registers: 8 LOP3 per adder instead of 3 XOR + 2, measured
// 42 instead of 26 LOP3 per 512-bit pair in the register-resident kernels.
#define CSA(h, l, a, b, c)                                                                      \
  {                                                                                             \
    uint32_t a_ = (a), b_ = (b), c_ = (c);                                                      \
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(l) : "r"(a_), "r"(b_), "r"(c_));                \
    asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(h) : "r"(a_), "r"(b_), "r"(c_));                \
  }

template <int PV>
__device__ __forceinline__ uint32_t popc16(const uint32_t (&x)[16]) {
  if (PV == 0) {
    uint32_t d = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) d += __popc(x[i]);
    return d;
  } else if (PV == 1) {
    uint32_t o[5], t[5];
#pragma unroll
    for (int g = 0; g < 5; ++g) CSA(t[g], o[g], x[3 * g], x[3 * g + 1], x[3 * g + 2]);
    uint32_t ones = __popc(o[0]) + __popc(o[1]) + __popc(o[2]) + __popc(o[3]) + __popc(o[4]) + __popc(x[15]);
    uint32_t twos = __popc(t[0]) + __popc(t[1]) + __popc(t[2]) + __popc(t[3]) + __popc(t[4]);
    return ones + 2 * twos;
  } else if (PV == 3) {
    // 7 adders: 9 POPC
    uint32_t o[7], t[7];
#pragma unroll
    for (int g = 0; g < 5; ++g) CSA(t[g], o[g], x[3 * g], x[3 * g + 1], x[3 * g + 2]);
    CSA(t[5], o[5], o[0], o[1], o[2]);
    CSA(t[6], o[6], o[3], o[4], x[15]);
    uint32_t ones = __popc(o[5]) + __popc(o[6]);
    uint32_t twos = __popc(t[0]) + __popc(t[1]) + __popc(t[2]) + __popc(t[3]) + __popc(t[4]) + __popc(t[5]) + __popc(t[6]);
    return ones + 2 * twos;
  } else if (PV == 4) {
    // 9 adders: 7 POPC
    uint32_t o[7], t[9], f[2];
#pragma unroll
    for (int g = 0; g < 5; ++g) CSA(t[g], o[g], x[3 * g], x[3 * g + 1], x[3 * g + 2]);
    CSA(t[5], o[5], o[0], o[1], o[2]);
    CSA(t[6], o[6], o[3], o[4], x[15]);
    CSA(f[0], t[7], t[0], t[1], t[2]);
    CSA(f[1], t[8], t[3], t[4], t[5]);
    uint32_t ones = __popc(o[5]) + __popc(o[6]);
    uint32_t twos = __popc(t[6]) + __popc(t[7]) + __popc(t[8]);
    uint32_t fours = __popc(f[0]) + __popc(f[1]);
    return ones + 2 * twos + 4 * fours;
  } else {
    // 11 adders: 5 POPC
    uint32_t o[7], t[10], f[4], e0;
#pragma unroll
    for (int g = 0; g < 5; ++g) CSA(t[g], o[g], x[3 * g], x[3 * g + 1], x[3 * g + 2]);
    CSA(t[5], o[5], o[0], o[1], o[2]);
    CSA(t[6], o[6], o[3], o[4], x[15]);
    CSA(f[0], t[7], t[0], t[1], t[2]);
    CSA(f[1], t[8], t[3], t[4], t[5]);
    CSA(f[2], t[9], t[6], t[7], t[8]);
    CSA(e0, f[3], f[0], f[1], f[2]);
    return __popc(o[5]) + __popc(o[6]) + 2 * __popc(t[9]) + 4 * __popc(f[3]) + 8 * __popc(e0);
  }
}

template <int PV>
__device__ __forceinline__ uint32_t popc8(const uint32_t (&x)[8]) {
  if (PV == 0) {
    uint32_t d = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) d += __popc(x[i]);
    return d;
  } else {
    uint32_t o0, o1, o2, t0, t1, t2, t3, f0;
    CSA(t0, o0, x[0], x[1], x[2]);
    CSA(t1, o1, x[3], x[4], x[5]);
    CSA(t2, o2, o0, o1, x[6]);
    CSA(f0, t3, t0, t1, t2);
    return __popc(o2) + __popc(x[7]) + 2 * __popc(t3) + 4 * __popc(f0);
  }
}

template <int W32>
__device__ __forceinline__ void load_query(const uint32_t* sq, uint32_t (&q)[W32]) {
  if (W32 == 2) {
    uint2 v = *reinterpret_cast<const uint2*>(sq);
    q[0] = v.x; q[1] = v.y;
  } else {
#pragma unroll
    for (int c = 0; c < W32 / 4; ++c) {
      uint4 v = reinterpret_cast<const uint4*>(sq)[c];
      q[4 * c + 0] = v.x; q[4 * c + 1] = v.y; q[4 * c + 2] = v.z; q[4 * c + 3] = v.w;
    }
  }
}

template <int W32, int PV>
__device__ __forceinline__ uint32_t hamming_rq(const uint32_t (&r)[W32], const uint32_t (&q)[W32]) {
  uint32_t x[W32];
#pragma unroll
  for (int i = 0; i < W32; ++i) x[i] = r[i] ^ q[i];
  if (W32 == 2) {
    return __popc(x[0]) + __popc(x[1]);
  } else if (W32 == 4) {
    if (PV == 0) return __popc(x[0]) + __popc(x[1]) + __popc(x[2]) + __popc(x[3]);
    uint32_t o, t;
    CSA(t, o, x[0], x[1], x[2]);
    return __popc(o) + __popc(x[3]) + 2 * __popc(t);
  } else if (W32 == 8) {
    uint32_t y[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) y[i] = x[i];
    return popc8<PV>(y);
  } else {
    uint32_t d = 0;
#pragma unroll
    for (int b = 0; b < W32 / 16; ++b) {
      uint32_t y[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) y[i] = x[16 * b + i];
      d += popc16<PV>(y);
    }
    return d;
  }
}

template <int W32, int PV>
__device__ __forceinline__ uint32_t hamming(const uint32_t (&r)[W32], const uint32_t* sq) {
  uint32_t q[W32];
  load_query<W32>(sq, q);
  return hamming_rq<W32, PV>(r, q);
}

// Hamming distance with the XORs of the first-level adder triples hoisted out of the pair loop.
// For a triple (x0, x1, x2), x_i = r_i ^ q_i, the adder's sum is (r0^r1^r2) ^ (q0^q1^q2): the two
// halves are computed once per row and once per query, so the sum costs one LOP3; with x0 and x1
// materialised the carry is a function of (x0, x1, sum) alone (x2 = sum ^ x0 ^ x1): four LOP3
// per adder instead of five, 25 instead of 30 per 512-bit pair.  The ALU pipe is what binds
// when many queries share a pass (80 % busy in the ncu capture of the 32-query kernel).
// sq: the query block [W32 words | tau | pad3 | per 16-word block: 5 triples, pad3]
#ifndef JBN_TRIPLES
#define JBN_TRIPLES 1
#endif
template <int W32>
__device__ __forceinline__ void row_triples(const uint32_t (&r)[W32], uint32_t (&rt)[(W32 / 16) * 5]) {
#pragma unroll
  for (int b = 0; b < W32 / 16; ++b)
#pragma unroll
    for (int g = 0; g < 5; ++g) rt[5 * b + g] = r[16 * b + 3 * g] ^ r[16 * b + 3 * g + 1] ^ r[16 * b + 3 * g + 2];
}
template <int W32>
__device__ __forceinline__ uint32_t hamming_t(const uint32_t (&r)[W32], const uint32_t (&rt)[(W32 / 16) * 5],
                                              const uint32_t* sq) {
  uint32_t q[W32];
  load_query<W32>(sq, q);
  uint32_t d = 0;
#pragma unroll
  for (int b = 0; b < W32 / 16; ++b) {
    const uint4 qa = *reinterpret_cast<const uint4*>(sq + W32 + 4 + 8 * b);
    const uint32_t q4 = sq[W32 + 4 + 8 * b + 4];
    const uint32_t qt[5] = {qa.x, qa.y, qa.z, qa.w, q4};
    uint32_t o[7], t[7];
#pragma unroll
    for (int g = 0; g < 5; ++g) {
      const uint32_t sum = rt[5 * b + g] ^ qt[g];
      const uint32_t x0 = r[16 * b + 3 * g] ^ q[16 * b + 3 * g];
      const uint32_t x1 = r[16 * b + 3 * g + 1] ^ q[16 * b + 3 * g + 1];
      uint32_t carry;   // majority(x0, x1, sum ^ x0 ^ x1)
      asm("lop3.b32 %0, %1, %2, %3, 0xD4;" : "=r"(carry) : "r"(x0), "r"(x1), "r"(sum));
      o[g] = sum; t[g] = carry;
    }
    const uint32_t x15 = r[16 * b + 15] ^ q[16 * b + 15];
    CSA(t[5], o[5], o[0], o[1], o[2]);
    CSA(t[6], o[6], o[3], o[4], x15);
    const uint32_t ones = __popc(o[5]) + __popc(o[6]);
    const uint32_t twos = __popc(t[0]) + __popc(t[1]) + __popc(t[2]) + __popc(t[3]) + __popc(t[4]) + __popc(t[5]) + __popc(t[6]);
    d += ones + 2 * twos;
  }
  return d;
}

// The same with two more adder levels (9 adders, 7 POPC per 512-bit pair instead of 9) for the
// popcount regime, where the XU pipe (POPC, a quarter of the ALU rate) and the ALU pipe (LOP3, adds,
// compares) are both near their limits: the weighted sum of the seven popcounts is accumulated with
// integer multiply-adds, which issue on the FMA pipe, so it costs neither of the two.  To keep the
// compiler from turning them back into adds and shifts the sum is kept scaled by HT4_SCALE (weights
// 7 / 14 / 28); the caller compares against the threshold scaled alike and divides on the rare path.
static const uint32_t HT4_SCALE = 7;
__device__ __forceinline__ uint32_t mad_u32(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
template <int W32>
__device__ __forceinline__ uint32_t hamming_t4(const uint32_t (&r)[W32], const uint32_t (&rt)[(W32 / 16) * 5],
                                               const uint32_t* sq) {
  uint32_t q[W32];
  load_query<W32>(sq, q);
  uint32_t acc = 0;
#pragma unroll
  for (int b = 0; b < W32 / 16; ++b) {
    const uint4 qa = *reinterpret_cast<const uint4*>(sq + W32 + 4 + 8 * b);
    const uint32_t q4 = sq[W32 + 4 + 8 * b + 4];
    const uint32_t qt[5] = {qa.x, qa.y, qa.z, qa.w, q4};
    uint32_t o[7], t[9], f[2];
#pragma unroll
    for (int g = 0; g < 5; ++g) {
      const uint32_t sum = rt[5 * b + g] ^ qt[g];
      const uint32_t x0 = r[16 * b + 3 * g] ^ q[16 * b + 3 * g];
      const uint32_t x1 = r[16 * b + 3 * g + 1] ^ q[16 * b + 3 * g + 1];
      uint32_t carry;
      asm("lop3.b32 %0, %1, %2, %3, 0xD4;" : "=r"(carry) : "r"(x0), "r"(x1), "r"(sum));
      o[g] = sum; t[g] = carry;
    }
    const uint32_t x15 = r[16 * b + 15] ^ q[16 * b + 15];
    CSA(t[5], o[5], o[0], o[1], o[2]);
    CSA(t[6], o[6], o[3], o[4], x15);
    CSA(f[0], t[7], t[0], t[1], t[2]);
    CSA(f[1], t[8], t[3], t[4], t[5]);
    acc = mad_u32(__popc(o[5]), HT4_SCALE, acc);
    acc = mad_u32(__popc(o[6]), HT4_SCALE, acc);
    acc = mad_u32(__popc(t[6]), 2 * HT4_SCALE, acc);
    acc = mad_u32(__popc(t[7]), 2 * HT4_SCALE, acc);
    acc = mad_u32(__popc(t[8]), 2 * HT4_SCALE, acc);
    acc = mad_u32(__popc(f[0]), 4 * HT4_SCALE, acc);
    acc = mad_u32(__popc(f[1]), 4 * HT4_SCALE, acc);
  }
  return acc;
}

// order-preserving image of a float (no NaN expected); -0 is folded into +0 first so that
// equal scores tie exactly like operator< on float does in the reference heap
__device__ __forceinline__ uint32_t float_key(float f) {
  f = __fadd_rn(f, 0.0f);
  uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ __forceinline__ float key_float(uint32_t k) {
  uint32_t b = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  float f; memcpy(&f, &b, 4); return f;
#endif
}

// largest key_hi that can still be below the exclusive bound `tau`
__host__ __device__ __forceinline__ uint32_t tau_hi_of(unsigned long long tau) {
  return tau ? (uint32_t)((tau - 1ull) >> 32) : 0u;
}

// euclid_lsh fast reject.  A row with norm n at Hamming distance d scores n * (n - c_d) with
// c_d = 2 * norm_q * cos(pi d / hash_num): a parabola in n.  It can reach the list only while its
// float score is <= T (the list's current bound), i.e. while n lies between the roots of
// n^2 - c_d n - T = 0.  The hot loop therefore needs no double arithmetic at all: one float
// lookup by d (the upper root) and one compare against the row norm; the exact score is computed
// only for the rare rows below it (rows under the lower root are rejected there as well, by the
// exact comparison).  The bound can only err towards admitting a row, and it is cheap enough to
// recompute whenever a list tightens (float arithmetic with directed rounding, no double sqrt):
//   T  -> two floats above it (anything below the next float may still round to T) plus
//         1.5e-14 * c^2, which covers the two double roundings of the exact score;
//   c  -> rounded up where it is added, away from zero where it is squared;
//   every operation rounds towards +inf, and the result is moved one more float up.
// A non-finite c or T disables the filter for that entry (+inf); a discriminant that is negative
// even after rounding up means no norm is admissible at that distance (-inf).
// upper root of n^2 - c n - Tu = 0, every operation rounded up (+inf: no filtering; -inf: nothing admissible)
__device__ __forceinline__ float euclid_upper_root(float c_add, float c_mag, float Tu) {
  const float disc = __fmaf_ru(c_mag, c_mag, __fmul_ru(4.0f, Tu));
  if (!(disc < INFINITY)) return INFINITY;                                // overflow or NaN: no filtering
  if (disc < 0.0f) return -INFINITY;
  const float hi = __fmul_ru(0.5f, __fadd_ru(c_add, __fsqrt_ru(disc)));
  if (!(fabsf(hi) < INFINITY)) return INFINITY;
  return __uint_as_float(hi >= 0.0f ? __float_as_uint(hi + 0.0f) + 1u : __float_as_uint(hi) - 1u);   // one more float up
}
// fmode: the score is computed in float (score_mode 1).  Its three roundings can move it by up to
// 2^-24 * n * (2 n + 3 |c|) (+ second-order terms) away from the exact value, n being the row norm; n is
// what the bound limits, so the slack is taken at a norm 1 % above a first estimate of the bound and the
// result is used only if it stays below that norm (else: no filtering at that distance).
__device__ __forceinline__ float euclid_bound_core(float c_add, float c_mag, uint32_t tau_hi, bool fmode = false) {
  const float Tf = key_float(tau_hi);
  if (!(fabsf(Tf) < INFINITY) || !(fabsf(c_mag) < INFINITY)) return INFINITY;
  float Tu = Tf + 0.0f;                                                   // -0 -> +0
  Tu = __uint_as_float(Tu >= 0.0f ? __float_as_uint(Tu) + 2u : (__float_as_uint(Tu) >= 0x80000002u ? __float_as_uint(Tu) - 2u : 1u));
  Tu = __fadd_ru(Tu, __fmaf_ru(1.5e-14f, __fmul_ru(c_mag, c_mag), 1e-37f));
  const float hi = euclid_upper_root(c_add, c_mag, Tu);
  if (!fmode || !(fabsf(hi) < INFINITY)) return hi;
  const float n1 = __fadd_ru(__fmul_ru(fabsf(hi), 1.01f), 1e-30f);
  const float slack = __fmul_ru(2.4e-7f, __fmul_ru(n1, __fadd_ru(__fmul_ru(2.0f, n1), __fmul_ru(3.0f, fabsf(c_mag)))));   // 4 * 2^-24: twice the first-order bound
  const float hi2 = euclid_upper_root(c_add, c_mag, __fadd_ru(Tu, slack));
  return (hi2 <= n1) ? hi2 : INFINITY;
}
__device__ __forceinline__ float euclid_bound(double c, uint32_t tau_hi, bool fmode = false) {
  if (!(fabs(c) < (double)INFINITY)) return INFINITY;
  const float c_add = __double2float_ru(c);
  const float c_mag = c >= 0.0 ? c_add : __double2float_rd(c);          // |c_mag| >= |c|
  return euclid_bound_core(c_add, c_mag, tau_hi, fmode);
}
// the same from c rounded up to float alone (the per-query table the TMA kernel keeps in shared memory,
// so that re-deriving a query's bounds touches neither global memory nor the double pipe): where c is
// squared, a negative c is taken one float further from zero, which covers the rounding of c_add
__device__ __forceinline__ float euclid_bound_f(float c_add, uint32_t tau_hi, bool fmode = false) {
  const float c_mag = c_add >= 0.0f ? c_add : __uint_as_float(__float_as_uint(c_add) + 1u);
  return euclid_bound_core(c_add, c_mag, tau_hi, fmode);
}

// ---------------------------------------------------------------------------
// per-(CTA, query) candidate list
// ---------------------------------------------------------------------------
// All list state lives in the kernel's dynamic shared memory.  The control block holds byte
// offsets from its base (not generic pointers), so every access below is an LDS/STS/ATOMS.
extern __shared__ __align__(128) unsigned char g_smem[];

struct ListCtl {
  uint32_t list_off;   // u64 [qp][cap]
  uint32_t tau_off;    // u64 [qp] exclusive upper bound on admissible keys
  uint32_t floor_off;  // u64 [qp] exclusive lower bound (pagination) when has_floor
  uint32_t cnt_off;    // u32 [qp]
  uint32_t lock_off;   // u32 [qp]
  uint32_t tauhi_off;  // u32: fast-reject word of query 0; query q at + q * tau_stride words
  uint32_t stash_off;  // per warp: u64 key[STASH] | u32 query[STASH] | u32 count | pad
  uint32_t tau_stride;
  uint32_t local_rows;
  uint32_t cap, k;
  uint32_t has_floor;
  uint32_t nq, q0, hist_stride;
  uint32_t* ghist;     // global: [nq] bound words | [nq][hist_stride] counts; nullptr when unused
  uint32_t no_flush_refresh;
  uint32_t score_mode;
  // euclid only
  uint32_t bnd_off;    // float [qp][hash_num + 1]: norm above which a row at that distance cannot be admitted
  uint32_t ctab_off;   // float [qp][hash_num + 1]: 2 * norm_q * cos(pi d / hash_num) rounded up, or NO_CTAB
  uint32_t cos_off, qn_off;   // with a ctab: double [hash_num + 1] cos table and float [qp] query norms in shared memory
  uint32_t hash_num;
  const double* coslut;   // global: cos(pi * d / hash_num)
  const float* qnorms;    // global: query norms
};
#ifndef JBN_REFRESH_MOD
#define JBN_REFRESH_MOD 4   // every JBN_REFRESH_MOD-th warp re-derives the shared bound after a flush
#endif
static const uint32_t NO_CTAB = 0xFFFFFFFFu;
__shared__ ListCtl s_L;   // written once by the prologue, read by the rare-path functions

__device__ __forceinline__ unsigned long long* sm_u64(uint32_t off) { return reinterpret_cast<unsigned long long*>(g_smem + off); }
__device__ __forceinline__ uint32_t* sm_u32(uint32_t off) { return reinterpret_cast<uint32_t*>(g_smem + off); }

// euclid: recompute query q's norm bounds for its current (tightened) threshold.  One full warp.
// Readers may see a mix of old and new entries: the bounds only ever shrink, so any mix still
// admits every admissible norm.
__device__ __noinline__ void euclid_bounds_update(uint32_t q, int lane) {
  const uint32_t th = *reinterpret_cast<volatile uint32_t*>(sm_u32(s_L.tauhi_off) + (size_t)q * s_L.tau_stride);
  float* b = reinterpret_cast<float*>(g_smem + s_L.bnd_off) + (size_t)q * (s_L.hash_num + 1);
  if (s_L.ctab_off != NO_CTAB) {
    const float* c = reinterpret_cast<const float*>(g_smem + s_L.ctab_off) + (size_t)q * (s_L.hash_num + 1);
    for (uint32_t d = lane; d <= s_L.hash_num; d += 32) b[d] = euclid_bound_f(c[d], th, s_L.score_mode != 0);
  } else {
    const double two_nq = (double)(2.0f * s_L.qnorms[s_L.q0 + q]);
    for (uint32_t d = lane; d <= s_L.hash_num; d += 32) b[d] = euclid_bound(__dmul_rn(two_nq, s_L.coslut[d]), th, s_L.score_mode != 0);
  }
  __syncwarp();
}

// Keep the k smallest of the cnt (distinct) keys of list q and tighten the threshold to the
// k-th smallest.  No sort: the k-th smallest key is found by a most-significant-bit-first search
// (count keys below a trial value, one warp reduction per bit), then the survivors are packed to
// the front in place.  Called by one full warp that holds lock q (or after the CTA-wide barrier
// at the end).  Returns the new count; the survivors are NOT ordered.
template <bool EUCLID>
__device__ __noinline__ uint32_t list_compact(uint32_t q, uint32_t cnt, int lane, bool last = false) {
  const uint32_t cap = s_L.cap, k = s_L.k;
  unsigned long long* a = sm_u64(s_L.list_off) + (size_t)q * cap;
  if (cnt <= k) return cnt;
  unsigned long long mx = 0;
  for (uint32_t i = lane; i < cnt; i += 32) mx = max(mx, a[i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(FULL_MASK, mx, o));
  unsigned long long prefix = 0;
  for (int bit = 63 - __clzll((long long)(mx | 1ull)); bit >= 0; --bit) {
    const unsigned long long trial = prefix | (1ull << bit);
    uint32_t c = 0;
    for (uint32_t i = lane; i < cnt; i += 32) c += (a[i] < trial) ? 1u : 0u;
    c = __reduce_add_sync(FULL_MASK, c);
    if (c < k) prefix = trial;   // fewer than k keys below trial: the k-th smallest is >= trial
  }
  // prefix is now the k-th smallest key; exactly k keys are <= prefix
  uint32_t w = 0;
  for (uint32_t base = 0; base < cnt; base += 32) {
    const uint32_t i = base + lane;
    const unsigned long long v = (i < cnt) ? a[i] : KEY_SENTINEL;
    const bool keep = v <= prefix;
    const unsigned m = __ballot_sync(FULL_MASK, keep);
    __syncwarp();
    if (keep) a[w + __popc(m & ((1u << lane) - 1u))] = v;
    w += __popc(m);
    __syncwarp();
  }
  if (lane == 0) {   // atomicMin: a tighter bound may have arrived from the shared histogram meanwhile
    atomicMin(sm_u64(s_L.tau_off) + q, prefix);
    atomicMin(sm_u32(s_L.tauhi_off) + (size_t)q * s_L.tau_stride, tau_hi_of(prefix));
  }
  __syncwarp();
  if (EUCLID && !last) euclid_bounds_update(q, lane);
  return w;
}

// Warp-uniform slow path: lanes with cand != 0 offer (key_hi, row).
template <bool EUCLID>
__device__ __noinline__ void list_offer(uint32_t q, uint32_t key_hi, uint32_t row, bool cand, int lane) {
  const unsigned long long key = ((unsigned long long)key_hi << 32) | row;
  volatile unsigned long long* tau_p = sm_u64(s_L.tau_off) + q;
  bool pass = cand && row < s_L.local_rows && key < *tau_p;  // rows of the padded last tile are not rows
  if (s_L.has_floor) pass = pass && key > sm_u64(s_L.floor_off)[q];
  if (!__any_sync(FULL_MASK, pass)) return;
  // per-query spin lock with acquire / release semantics on the atomics themselves (no
  // stand-alone fences: the critical section is a handful of shared-memory accesses)
  const uint32_t lock = (uint32_t)__cvta_generic_to_shared(sm_u32(s_L.lock_off) + q);
  if (lane == 0) {
    uint32_t old;
    do {
      asm volatile("atom.acquire.cta.shared.cas.b32 %0, [%1], 0, 1;" : "=r"(old) : "r"(lock) : "memory");
    } while (old != 0u);
  }
  __syncwarp();
  pass = pass && key < *tau_p;  // threshold may have tightened while we waited
  const unsigned m = __ballot_sync(FULL_MASK, pass);
  if (m) {
    volatile uint32_t* cnt_p = sm_u32(s_L.cnt_off) + q;
    uint32_t cnt = *cnt_p;
    if (pass) {
      uint32_t pos = cnt + __popc(m & ((1u << lane) - 1u));
      sm_u64(s_L.list_off)[(size_t)q * s_L.cap + pos] = key;
    }
    cnt += __popc(m);
    __syncwarp();
    if (cnt > s_L.cap - 32) cnt = list_compact<EUCLID>(q, cnt, lane);
    if (lane == 0) *cnt_p = cnt;
  }
  __syncwarp();
  if (lane == 0) {
    uint32_t old;
    asm volatile("atom.release.cta.shared.exch.b32 %0, [%1], 0;" : "=r"(old) : "r"(lock) : "memory");
  }
}

// Every inserted candidate is also counted in a histogram shared by all CTAs (global memory).
// If the counts of the bins up to B add up to k, k rows with a key inside those bins exist, so
// no key above bin B can be in the top k: a valid bound for every CTA, and unlike a CTA's own
// k-th best it reflects what the whole GPU has seen.
//   hamming engines: one bin per distance (key_hi = distance).
//   euclid_lsh: key_hi is the order-preserving image of the float score; its upper 16 bits (sign,
//   exponent, 7 mantissa bits: steps of under 1 % of the score) index 65536 fine bins, summarised
//   by 256 coarse bins (upper 8 bits) so that a refresh reads two short runs instead of a long one.
static const uint32_t EUCLID_COARSE_BINS = 256, EUCLID_FINE_BINS = 65536;
__device__ __forceinline__ void apply_shared_bound(uint32_t q, uint32_t key_hi) {
  atomicMin(sm_u64(s_L.tau_off) + q, ((unsigned long long)key_hi + 1ull) << 32);
  atomicMin(sm_u32(s_L.tauhi_off) + (size_t)q * s_L.tau_stride, key_hi);
}
template <bool EUCLID>
__device__ __forceinline__ void hist_count(uint32_t gq, uint32_t key_hi) {
  uint32_t* h = s_L.ghist + s_L.nq + (size_t)gq * s_L.hist_stride;
  if (EUCLID) {
    atomicAdd(h + EUCLID_COARSE_BINS + (key_hi >> 16), 1u);   // fine first, then its coarse summary
    atomicAdd(h + (key_hi >> 24), 1u);
  } else {
    atomicAdd(h + key_hi, 1u);
  }
}

template <bool EUCLID>
__device__ __noinline__ void hist_refresh(uint32_t q, int lane);

// hamming engines: one bin per distance
template <>
__device__ __noinline__ void hist_refresh<false>(uint32_t q, int lane) {
  const uint32_t gq = s_L.q0 + q;
  const volatile uint32_t* h = s_L.ghist + s_L.nq + (size_t)gq * s_L.hist_stride;
  const uint32_t cur = *reinterpret_cast<volatile uint32_t*>(sm_u32(s_L.tauhi_off) + (size_t)q * s_L.tau_stride);
  const uint32_t limit = min(cur, s_L.hist_stride - 1);   // only a bound below the current one is of interest
  const uint32_t k = s_L.k;
  uint32_t run = 0;
  // eight 32-bin blocks per round, their loads all in flight together (one L2 round trip per
  // 256 distances instead of one per 32)
  for (uint32_t base0 = 0; base0 <= limit; base0 += 256) {
    uint32_t v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const uint32_t d = base0 + 32u * j + lane;
      v[j] = (d <= limit) ? h[d] : 0u;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const uint32_t tot = __reduce_add_sync(FULL_MASK, v[j]);
      if (run + tot >= k) {   // the crossing is inside this block: locate it with one prefix scan
        uint32_t x = v[j];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t n = __shfl_up_sync(FULL_MASK, x, o);
          if (lane >= o) x += n;
        }
        const unsigned hit = __ballot_sync(FULL_MASK, run + x >= k);
        const uint32_t D = base0 + 32u * j + (__ffs(hit) - 1);
        if (lane == 0 && D < cur) {
          apply_shared_bound(q, D);
          atomicMin(s_L.ghist + gq, D);   // publish for CTAs that insert rarely
        }
        __syncwarp();
        return;
      }
      run += tot;
    }
  }
  __syncwarp();
}

// one 256-bin block [base0, base0 + 256) clipped at `last`: all eight loads of a lane in flight together
__device__ __forceinline__ void hist_load8(const volatile uint32_t* h, uint32_t base0, uint32_t last, int lane, uint32_t (&v)[8]) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint32_t d = base0 + 32u * j + lane;
    v[j] = (d <= last) ? h[d] : 0u;
  }
}
// first bin of the loaded block at which run + (counts up to it) reaches k, or 0xFFFFFFFF; `run` is
// advanced to the count of the bins before the returned one (or past the block)
__device__ __forceinline__ uint32_t hist_scan8(const uint32_t (&v)[8], uint32_t base0, uint32_t& run, uint32_t k, int lane) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint32_t tot = __reduce_add_sync(FULL_MASK, v[j]);
    if (run + tot >= k) {
      uint32_t x = v[j];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(FULL_MASK, x, o);
        if (lane >= o) x += n;
      }
      const unsigned hit = __ballot_sync(FULL_MASK, run + x >= k);
      const int l = __ffs(hit) - 1;
      run += __shfl_sync(FULL_MASK, x - v[j], l);
      return base0 + 32u * j + l;
    }
    run += tot;
  }
  return 0xFFFFFFFFu;
}

// euclid_lsh: coarse bins locate the 256 fine bins in which the count reaches k.  The bound rarely
// leaves its coarse bin between two refreshes, so the fine bins under the current bound's coarse bin
// are fetched together with the coarse ones: one L2 round trip instead of two in the common case.
template <>
__device__ __noinline__ void hist_refresh<true>(uint32_t q, int lane) {
  const uint32_t gq = s_L.q0 + q;
  const volatile uint32_t* h = s_L.ghist + s_L.nq + (size_t)gq * s_L.hist_stride;
  const uint32_t cur = *reinterpret_cast<volatile uint32_t*>(sm_u32(s_L.tauhi_off) + (size_t)q * s_L.tau_stride);
  const uint32_t last16 = cur >> 16, Cg = last16 >> 8;
  uint32_t vc[8], vf[8];
  hist_load8(h, 0, Cg, lane, vc);
  hist_load8(h + EUCLID_COARSE_BINS, Cg << 8, last16, lane, vf);
  uint32_t run = 0;
  const uint32_t C = hist_scan8(vc, 0, run, s_L.k, lane);
  if (C == 0xFFFFFFFFu) return;
  if (C != Cg) hist_load8(h + EUCLID_COARSE_BINS, C << 8, (C << 8) | 0xFFu, lane, vf);   // (C < Cg: the whole coarse bin lies below the bound)
  const uint32_t word = hist_scan8(vf, C << 8, run, s_L.k, lane);
  if (word == 0xFFFFFFFFu) return;
  const uint32_t bound = (word << 16) | 0xFFFFu;   // every key of that fine bin stays admissible
  if (bound < cur) {
    if (lane == 0) {
      apply_shared_bound(q, bound);
      atomicMin(s_L.ghist + gq, word);   // publish for CTAs that insert rarely
    }
    __syncwarp();
    euclid_bounds_update(q, lane);
  }
  __syncwarp();
}

// Insert everything this warp has stashed (grouped by query, one list_offer per distinct query).
template <bool EUCLID>
__device__ __noinline__ void stash_flush(int warp, int lane, bool refresh = true) {
  unsigned char* base = g_smem + s_L.stash_off + (size_t)warp * (32 * 8 + 32 * 4 + 16);
  unsigned long long* skey = reinterpret_cast<unsigned long long*>(base);
  uint32_t* sq = reinterpret_cast<uint32_t*>(base + 32 * 8);
  volatile uint32_t* scnt = reinterpret_cast<uint32_t*>(base + 32 * 8 + 32 * 4);
  const uint32_t n = *scnt;
  __syncwarp();
  const bool have = (uint32_t)lane < n;
  const unsigned long long key = have ? skey[lane] : 0ull;
  const uint32_t q = have ? sq[lane] : 0xFFFFFFFFu;
  __syncwarp();
  if (lane == 0) *scnt = 0;
  unsigned todo = __ballot_sync(FULL_MASK, have);
  while (todo) {
    const uint32_t qcur = __shfl_sync(FULL_MASK, q, __ffs(todo) - 1);
    const bool sel = have && q == qcur;
    list_offer<EUCLID>(qcur, (uint32_t)(key >> 32), (uint32_t)key, sel, lane);
    if (refresh && s_L.ghist && (warp % JBN_REFRESH_MOD) == 0 && !s_L.no_flush_refresh) hist_refresh<EUCLID>(qcur, lane);
    todo &= ~__ballot_sync(FULL_MASK, sel);
  }
}

// One warp per CTA polls the bounds other CTAs have published (a coalesced read of nqp words).
template <bool EUCLID>
__device__ __forceinline__ void poll_shared_bounds(const ListCtl& L, uint32_t nqp, int lane) {
  if (L.ghist == nullptr) return;
  if (!EUCLID) {
    for (uint32_t q = lane; q < nqp; q += 32) {
      const uint32_t D = *reinterpret_cast<const volatile uint32_t*>(L.ghist + L.q0 + q);
      if (D < *reinterpret_cast<const volatile uint32_t*>(sm_u32(L.tauhi_off) + (size_t)q * L.tau_stride)) {
        atomicMin(sm_u64(L.tau_off) + q, ((unsigned long long)D + 1ull) << 32);
        atomicMin(sm_u32(L.tauhi_off) + (size_t)q * L.tau_stride, D);
      }
    }
    return;
  }
  for (uint32_t qb = 0; qb < nqp; qb += 32) {
    const uint32_t q = qb + lane;
    bool tighter = false;
    if (q < nqp) {
      const uint32_t word = *reinterpret_cast<const volatile uint32_t*>(L.ghist + L.q0 + q);
      const uint32_t bound = (word << 16) | 0xFFFFu;
      tighter = bound < *reinterpret_cast<const volatile uint32_t*>(sm_u32(L.tauhi_off) + (size_t)q * L.tau_stride);
      if (tighter) {
        atomicMin(sm_u64(L.tau_off) + q, ((unsigned long long)bound + 1ull) << 32);
        atomicMin(sm_u32(L.tauhi_off) + (size_t)q * L.tau_stride, bound);
      }
    }
    {   // the norm bounds of the tightened queries follow
      unsigned m = __ballot_sync(FULL_MASK, tighter);
      while (m) {
        const int l = __ffs(m) - 1;
        m &= m - 1;
        euclid_bounds_update(qb + l, lane);
      }
    }
  }
}

// Warp-uniform: lanes with cand offer (key_hi, row) for query q; admissible ones are stashed.
template <bool EUCLID>
__device__ __noinline__ void stash_push(uint32_t q, uint32_t key_hi, uint32_t row, bool cand, int warp, int lane) {
  const unsigned long long key = ((unsigned long long)key_hi << 32) | row;
  bool pass = cand && row < s_L.local_rows &&
              key < *reinterpret_cast<volatile unsigned long long*>(sm_u64(s_L.tau_off) + q);
  if (s_L.has_floor) pass = pass && key > sm_u64(s_L.floor_off)[q];
  const unsigned m = __ballot_sync(FULL_MASK, pass);
  if (!m) return;
  // counted in the GPU-wide histogram right away, not when the stash is flushed: at the start of a pass
  // every CTA holds a few candidates per warp, and the bound all of them together imply is what ends
  // that phase (each (row, query) pair comes by here once, so no row is counted twice)
  if (s_L.ghist && pass) hist_count<EUCLID>(s_L.q0 + q, key_hi);
  unsigned char* base = g_smem + s_L.stash_off + (size_t)warp * (32 * 8 + 32 * 4 + 16);
  volatile uint32_t* scnt = reinterpret_cast<uint32_t*>(base + 32 * 8 + 32 * 4);
  uint32_t n = *scnt;
  if (n + __popc(m) > 32) {   // no room: insert what is stashed now
    stash_flush<EUCLID>(warp, lane);
    n = 0;
  }
  if (pass) {
    const uint32_t pos = n + __popc(m & ((1u << lane) - 1u));
    reinterpret_cast<unsigned long long*>(base)[pos] = key;
    reinterpret_cast<uint32_t*>(base + 32 * 8)[pos] = q;
  }
  __syncwarp();
  if (lane == 0) *scnt = n + __popc(m);
  __syncwarp();
}

// the engine's score of a row: n_i * (n_i - 2 n_q cos(theta_d)).  mode 0 (frozen choice): the expression in
// double, one rounding per operation, narrowed to float; mode 1 (SURVEY A.9 alternate: std::cos on a float
// picks cosf, the expression stays float): the table holds cosf values and every operation is a float one
__device__ __forceinline__ float euclid_exact_score(float nif, float two_nq, double cosd, uint32_t mode) {
  if (mode) return __fmul_rn(nif, __fsub_rn(nif, __fmul_rn(two_nq, (float)cosd)));
  const double ni = (double)nif;
  return __double2float_rn(__dmul_rn(ni, __dsub_rn(ni, __dmul_rn((double)two_nq, cosd))));
}

// euclid slow path, warp-uniform: lanes with cand compute the exact score of (row, query q) --
// double arithmetic with one rounding per operation, narrowed to float: the same values as the
// CPU engine -- and offer it.
__device__ __noinline__ void euclid_offer(uint32_t q, uint32_t d, float nif, uint32_t row, bool cand, int warp, int lane) {
  if (!__any_sync(FULL_MASK, cand)) return;
  uint32_t key = 0;
  if (cand) {
    float two_nq;
    double cosd;
    if (s_L.ctab_off != NO_CTAB) {   // TMA kernel: both operands sit in shared memory
      two_nq = 2.0f * reinterpret_cast<const float*>(g_smem + s_L.qn_off)[q];
      cosd = reinterpret_cast<const double*>(g_smem + s_L.cos_off)[d];
    } else {
      two_nq = 2.0f * s_L.qnorms[s_L.q0 + q];
      cosd = s_L.coslut[d];
    }
    key = float_key(euclid_exact_score(nif, two_nq, cosd, s_L.score_mode));
  }
  stash_push<true>(q, key, row, cand, warp, lane);
}
__device__ __forceinline__ bool norm_inside(float n, float hi) { return !(n > hi); }   // NaN norms go to the exact path

__device__ __forceinline__ uint32_t stash_count(const ListCtl& L, int warp) {
  return *reinterpret_cast<const volatile uint32_t*>(g_smem + L.stash_off + (size_t)warp * (32 * 8 + 32 * 4 + 16) + 32 * 8 + 32 * 4);
}
// Insertions are batched: one lock acquisition moves up to 32 candidates, which matters while a
// CTA's threshold is still loose and every tile yields some.
static const uint32_t STASH_FLUSH_AT = 16;
static const uint32_t EARLY_REFRESH_STAGES = 4;

__host__ __device__ inline size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }
// a query's shared-memory block: W32 signature words, then the fast-reject threshold word
// (tau_hi), padded to 16 bytes -- one base address serves the whole inner loop
// ... then, for rows of 512 bits and more, the XOR of each of the query's first-level adder
// triples (5 words per 16-word block, padded to 8): see hamming_t
__host__ __device__ constexpr uint32_t q_stride(uint32_t w32) { return w32 + 4 + (w32 >= 16 ? (w32 / 16) * 8 : 0); }

// per-warp stash of candidates whose list insertion is deferred until the warp has released
// the stage it was reading (so a slow insertion never holds up the memory pipeline)
static const uint32_t STASH = 32;                       // entries per warp
static const uint32_t STASH_BYTES = STASH * 8 + STASH * 4 + 16;
static const uint32_t STASH_WARPS = 33;

struct SmemPlan {
  size_t q_off, lut_off, ctab_off, cos_off, qn_off, list_off, tau_off, floor_off, cnt_off, lock_off, tauhi_off, stash_off, total;
};
__host__ __device__ inline SmemPlan plan_smem(uint32_t qp, uint32_t w32, uint32_t cap, bool euclid,
                                              uint32_t hash_num, size_t base, bool ctab = false) {
  SmemPlan p;
  size_t o = base;
  p.q_off = o; o = align16(o + (size_t)qp * q_stride(w32) * 4);
  p.lut_off = o; if (euclid) o = align16(o + (size_t)qp * (hash_num + 1) * 4);
  p.ctab_off = o; if (euclid && ctab) o = align16(o + (size_t)qp * (hash_num + 1) * 4);
  p.cos_off = o; if (euclid && ctab) o = align16(o + (size_t)(hash_num + 1) * 8);   // cos(pi d / hash_num), once per CTA
  p.qn_off = o; if (euclid && ctab) o = align16(o + (size_t)qp * 4);                // the group's query norms
  p.list_off = o; o = align16(o + (size_t)qp * cap * 8);
  p.tau_off = o; o = align16(o + (size_t)qp * 8);
  p.floor_off = o; o = align16(o + (size_t)qp * 8);
  p.cnt_off = o; o = align16(o + (size_t)qp * 4);
  p.lock_off = o; o = align16(o + (size_t)qp * 4);
  p.tauhi_off = o;  // (tau_hi lives inside the query blocks)
  p.stash_off = o; o = align16(o + (size_t)STASH_WARPS * STASH_BYTES);
  p.total = o;
  return p;
}

// ---------------------------------------------------------------------------
// one tile (lane = row) against the CTA's queries
// ---------------------------------------------------------------------------
template <int W32, bool EUCLID, int PV>
__device__ __forceinline__ void rank_tile(const uint32_t (&cur)[W32], uint32_t row, float ni, uint32_t nqp,
                                          const uint32_t* s_q, const float* s_bnd, uint32_t hash_num,
                                          const ListCtl& L, int lane) {
  constexpr uint32_t QS = q_stride(W32);
  constexpr bool TRIPLES = JBN_TRIPLES && W32 >= 16 && (PV == 3 || PV == 4);
  constexpr bool SCALED = TRIPLES && PV == 4 && !EUCLID;   // distances carried as HT4_SCALE * d (hamming_t4)
  constexpr int NT = W32 >= 16 ? (W32 / 16) * 5 : 1;
  const int warp = threadIdx.x >> 5;
  uint32_t rt[NT];
  if constexpr (TRIPLES) row_triples<W32>(cur, rt);
  uint32_t q = 0;
  for (; q + 2 <= nqp; q += 2) {  // two independent distance chains in flight
    const uint32_t* b0 = s_q + q * QS;
    const uint32_t* b1 = b0 + QS;
    uint32_t d0, d1;
    if constexpr (SCALED) {
      d0 = hamming_t4<W32>(cur, rt, b0);
      d1 = hamming_t4<W32>(cur, rt, b1);
    } else if constexpr (TRIPLES) {
      d0 = hamming_t<W32>(cur, rt, b0);
      d1 = hamming_t<W32>(cur, rt, b1);
    } else {
      d0 = hamming<W32, PV>(cur, b0);
      d1 = hamming<W32, PV>(cur, b1);
    }
    if (EUCLID) {
      const bool c0 = norm_inside(ni, s_bnd[q * (hash_num + 1) + d0]);
      const bool c1 = norm_inside(ni, s_bnd[(q + 1) * (hash_num + 1) + d1]);
      if (__any_sync(FULL_MASK, c0 || c1)) {
        euclid_offer(q, d0, ni, row, c0, warp, lane);
        euclid_offer(q + 1, d1, ni, row, c1, warp, lane);
      }
    } else {
      const uint32_t t0 = *reinterpret_cast<const volatile uint32_t*>(b0 + W32);
      const uint32_t t1 = *reinterpret_cast<const volatile uint32_t*>(b1 + W32);
      // (a threshold of 0xFFFFFFFF, "none yet", must not wrap when scaled)
      const bool c0 = SCALED ? (d0 <= mad_u32(min(t0, 0x10000u), HT4_SCALE, 0u)) : (d0 <= t0);
      const bool c1 = SCALED ? (d1 <= mad_u32(min(t1, 0x10000u), HT4_SCALE, 0u)) : (d1 <= t1);
      if (__any_sync(FULL_MASK, c0 || c1)) {
        stash_push<false>(q, SCALED ? d0 / HT4_SCALE : d0, row, c0, warp, lane);
        stash_push<false>(q + 1, SCALED ? d1 / HT4_SCALE : d1, row, c1, warp, lane);
      }
    }
  }
  if (q < nqp) {
    const uint32_t* b0 = s_q + q * QS;
    uint32_t d0;
    if constexpr (TRIPLES) d0 = hamming_t<W32>(cur, rt, b0); else d0 = hamming<W32, PV>(cur, b0);
    if (EUCLID) {
      const bool c0 = norm_inside(ni, s_bnd[q * (hash_num + 1) + d0]);
      if (__any_sync(FULL_MASK, c0)) euclid_offer(q, d0, ni, row, c0, warp, lane);
    } else {
      const bool c0 = d0 <= *reinterpret_cast<const volatile uint32_t*>(b0 + W32);
      if (__any_sync(FULL_MASK, c0)) stash_push<false>(q, d0, row, c0, warp, lane);
    }
  }
}

// two tiles at once (lane holds row A and row B): every query word fetched from shared memory
// serves two rows, and the two distance chains are independent
template <int W32, bool EUCLID, int PV>
__device__ __forceinline__ void rank_tile2(const uint32_t (&ra)[W32], const uint32_t (&rb)[W32], uint32_t rowa,
                                           uint32_t rowb, float nia, float nib, uint32_t nqp, const uint32_t* s_q,
                                           const float* s_bnd, uint32_t hash_num, const ListCtl& L, int lane) {
  constexpr uint32_t QS = q_stride(W32);
  const int warp = threadIdx.x >> 5;
  const uint32_t* b0 = s_q;
#pragma unroll 2
  for (uint32_t q = 0; q < nqp; ++q, b0 += QS) {
    uint32_t qw[W32];
    load_query<W32>(b0, qw);
    const uint32_t da = hamming_rq<W32, PV>(ra, qw);
    const uint32_t db = hamming_rq<W32, PV>(rb, qw);
    if (EUCLID) {
      const bool ca = norm_inside(nia, s_bnd[q * (hash_num + 1) + da]);
      const bool cb = norm_inside(nib, s_bnd[q * (hash_num + 1) + db]);
      if (__any_sync(FULL_MASK, ca || cb)) {
        euclid_offer(q, da, nia, rowa, ca, warp, lane);
        euclid_offer(q, db, nib, rowb, cb, warp, lane);
      }
    } else {
      const uint32_t tau = *reinterpret_cast<const volatile uint32_t*>(b0 + W32);
      const bool ca = da <= tau, cb = db <= tau;
      if (__any_sync(FULL_MASK, ca || cb)) {
        stash_push<false>(q, da, rowa, ca, warp, lane);
        stash_push<false>(q, db, rowb, cb, warp, lane);
      }
    }
  }
}

// Register-resident queries (TMA variant, <= 4 queries per pass): the CTA's queries are loaded
// into registers once, thresholds are re-read once per stage, so the inner loop touches shared
// memory only for the rows themselves (and, for euclid, the norm intervals).  NR rows (tiles)
// are ranked together.
template <int W32, bool EUCLID, int PV, int NQ, int NR>
__device__ __forceinline__ void rank_reg(const uint32_t (&rows)[NR][W32], const uint32_t (&rowid)[NR],
                                         const float (&ni)[NR], const uint32_t (&qreg)[NQ][W32],
                                         const uint32_t (&tau)[NQ], const float* s_bnd, uint32_t hash_num,
                                         int warp, int lane) {
  uint32_t dist[NQ][NR];
  bool any = false;
#pragma unroll
  for (int q = 0; q < NQ; ++q) {
#pragma unroll
    for (int r = 0; r < NR; ++r) {
      const uint32_t d = hamming_rq<W32, PV>(rows[r], qreg[q]);
      dist[q][r] = d;
      any = any || (EUCLID ? norm_inside(ni[r], s_bnd[q * (hash_num + 1) + d]) : (d <= tau[q]));
    }
  }
  if (__any_sync(FULL_MASK, any)) {
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
#pragma unroll
      for (int r = 0; r < NR; ++r) {
        if (EUCLID)
          euclid_offer(q, dist[q][r], ni[r], rowid[r], norm_inside(ni[r], s_bnd[q * (hash_num + 1) + dist[q][r]]), warp, lane);
        else
          stash_push<false>(q, dist[q][r], rowid[r], dist[q][r] <= tau[q], warp, lane);
      }
    }
  }
}

// shared prologue: query blocks, euclid norm intervals, list control
template <int W32, bool EUCLID, bool CTAB = false>
__device__ __forceinline__ void scan_prologue(const ScanArgs& a, unsigned char* smem, const SmemPlan& sp, ListCtl& L,
                                              uint32_t*& s_q, float*& s_bnd, uint32_t q0, uint32_t nqp, int tid,
                                              int nthreads) {
  constexpr uint32_t QS = q_stride(W32);
  s_q = reinterpret_cast<uint32_t*>(smem + sp.q_off);
  s_bnd = reinterpret_cast<float*>(smem + sp.lut_off);
  L.list_off = (uint32_t)sp.list_off; L.tau_off = (uint32_t)sp.tau_off; L.floor_off = (uint32_t)sp.floor_off;
  L.cnt_off = (uint32_t)sp.cnt_off; L.lock_off = (uint32_t)sp.lock_off;
  L.tauhi_off = (uint32_t)(sp.q_off + W32 * 4);
  L.stash_off = (uint32_t)sp.stash_off;
  L.tau_stride = QS;
  L.local_rows = (uint32_t)a.local_rows;
  L.cap = a.cap; L.k = a.k; L.has_floor = (a.floor_keys != nullptr) ? 1u : 0u;
  L.nq = a.nq; L.q0 = q0; L.hist_stride = a.hist_stride; L.ghist = a.ghist;
  L.no_flush_refresh = (a.debug & 8u) ? 1u : 0u;
  L.score_mode = a.score_mode;
  L.bnd_off = (uint32_t)sp.lut_off; L.hash_num = a.hash_num;
  L.ctab_off = (EUCLID && CTAB) ? (uint32_t)sp.ctab_off : NO_CTAB;
  L.cos_off = (uint32_t)sp.cos_off; L.qn_off = (uint32_t)sp.qn_off;
  if (EUCLID && CTAB && (uint32_t)tid < nqp) reinterpret_cast<float*>(smem + sp.qn_off)[tid] = a.qnorms[q0 + tid];
  L.coslut = EUCLID ? a.coslut : nullptr; L.qnorms = EUCLID ? a.qnorms : nullptr;
  if (tid == 0) s_L = L;
  if ((uint32_t)tid < STASH_WARPS) *reinterpret_cast<uint32_t*>(smem + sp.stash_off + (size_t)tid * STASH_BYTES + STASH * 12) = 0u;
  for (uint32_t i = tid; i < nqp * QS; i += nthreads) {
    uint32_t q = i / QS, w = i % QS;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(a.qsigs + (size_t)(q0 + q) * a.w64);
    uint32_t v;
    if (w < W32) {
      v = (w < 2 * a.w64) ? (src[w] & sig_word_mask(a.hash_num, w)) : 0u;   // bits at or above hash_num never count
    } else if (w < W32 + 4) {
      v = tau_hi_of(a.tau_init ? a.tau_init[q0 + q] : KEY_SENTINEL);   // word W32; the pad words carry the same
    } else {   // XOR of the adder triple g of 16-word block b (hamming_t)
      const uint32_t b = (w - W32 - 4) / 8, g = (w - W32 - 4) % 8;
      v = 0u;
      if (g < 5)
        for (uint32_t j = 0; j < 3; ++j) {
          const uint32_t ww = 16 * b + 3 * g + j;
          if (ww < 2 * a.w64) v ^= src[ww] & sig_word_mask(a.hash_num, ww);
        }
    }
    s_q[i] = v;
  }
  if (EUCLID) {
    const uint32_t hn = a.hash_num + 1;
    for (uint32_t i = tid; i < nqp * hn; i += nthreads) {
      uint32_t q = i / hn, h = i % hn;
      float two_nq = 2.0f * a.qnorms[q0 + q];
      const double c = __dmul_rn((double)two_nq, a.coslut[h]);
      const uint32_t th = tau_hi_of(a.tau_init ? a.tau_init[q0 + q] : KEY_SENTINEL);
      if (CTAB) {
        const float c_add = (fabs(c) < (double)INFINITY) ? __double2float_ru(c) : INFINITY;
        reinterpret_cast<float*>(smem + sp.ctab_off)[i] = c_add;
        s_bnd[i] = euclid_bound_f(c_add, th, a.score_mode != 0);
      } else {
        s_bnd[i] = euclid_bound(c, th, a.score_mode != 0);
      }
    }
  }
  if ((uint32_t)tid < nqp) {
    sm_u32(L.cnt_off)[tid] = 0; sm_u32(L.lock_off)[tid] = 0;
    sm_u64(L.tau_off)[tid] = a.tau_init ? a.tau_init[q0 + tid] : KEY_SENTINEL;
    sm_u64(L.floor_off)[tid] = a.floor_keys ? a.floor_keys[q0 + tid] : 0ull;
  }
}

// ---------------------------------------------------------------------------
// variant 1: rows streamed with coalesced 128-bit LDG, one tile prefetched per warp
// ---------------------------------------------------------------------------
template <int W32, bool EUCLID, int PV>
#ifndef JBN_V1_PAIR
#define JBN_V1_PAIR 1
#endif
#ifndef JBN_V1_PAIR_MAX_W32
#define JBN_V1_PAIR_MAX_W32 4
#endif
// rows of 512 bits and more: one CTA per SM with 128 registers per thread (the row, its
// prefetched successor, the adder triples and two queries fit without spilling: +16 % at 32
// queries per pass, +17..42 % on 1024-bit rows); shorter rows: two CTAs per SM at 64 registers
// (measured 13..33 % faster there than one)
__global__ void __launch_bounds__(SCAN_THREADS, W32 >= 16 ? 1 : 2) scan_ldg_kernel(ScanArgs a) {
  unsigned char* smem = g_smem;
  const SmemPlan sp = plan_smem(a.qp, W32, a.cap, EUCLID, a.hash_num, 0);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t q0 = blockIdx.y * a.qp;
  const uint32_t nqp = min(a.qp, a.nq - q0);
  ListCtl L;
  uint32_t* s_q;
  float* s_bnd;
  scan_prologue<W32, EUCLID>(a, smem, sp, L, s_q, s_bnd, q0, nqp, tid, SCAN_THREADS);
  __syncthreads();

  // rows of <= 128 bits: a warp takes two adjacent tiles per step, so that every query word
  // fetched from shared memory serves two rows (+6..11 % measured; at 256 bits the second row
  // costs more in spilled registers than it saves: 6.0 k against 6.4 k queries/s at 24 per pass)
  constexpr bool PAIR = JBN_V1_PAIR && W32 <= JBN_V1_PAIR_MAX_W32;
  constexpr uint64_t TPS = PAIR ? 2 : 1;   // tiles per warp and step
  uint64_t t, t_end, t_step;
  if (a.interleave) {
    t = ((uint64_t)blockIdx.x * SCAN_WARPS + warp) * TPS;
    t_end = a.ntiles;
    t_step = (uint64_t)gridDim.x * SCAN_WARPS * TPS;
  } else {
    const uint64_t t_begin = (uint64_t)blockIdx.x * a.tiles_per_chunk;
    t = t_begin + warp * TPS;
    t_end = min(a.ntiles, t_begin + (uint64_t)a.tiles_per_chunk);
    t_step = SCAN_WARPS * TPS;
  }
  uint32_t poll_tick = 0;
  // Rows of 512 bits and more (one CTA per SM, 128 registers): the next tile is loaded into registers
  // while the current one is ranked.  Shorter rows run two CTAs per SM at 64 registers, where a row
  // held across the calls of the insertion path was kept on the stack -- the store sat right behind
  // its load and waited for it (6 % of all samples in the ncu capture of the 128-bit kernel).  There
  // the next tile is only pulled into L2 and loaded when its turn comes: 32 resident warps per SM
  // hide an L2 hit.
  constexpr bool REG_PREFETCH = W32 >= 16;
  if constexpr (PAIR) {
    uint32_t ca[W32], cb[W32];
    float fa = 0.0f, fb = 0.0f;
    for (; t < t_end; t += t_step) {
      const bool two = t + 1 < t_end;
      load_row<W32>(a.slab, t, lane, ca);
      if (EUCLID) fa = ld_stream_f32(a.norms + t * 32 + lane);
      if (two) {
        load_row<W32>(a.slab, t + 1, lane, cb);
        if (EUCLID) fb = ld_stream_f32(a.norms + (t + 1) * 32 + lane);
      }
      const uint64_t tn = t + t_step;
      if (tn < t_end) {
        prefetch_row_l2<W32>(a.slab, tn, lane);
        if (tn + 1 < t_end) prefetch_row_l2<W32>(a.slab, tn + 1, lane);
      }
      if (two)
        rank_tile2<W32, EUCLID, PV>(ca, cb, (uint32_t)(t * 32 + lane), (uint32_t)(t * 32 + 32 + lane), fa, fb, nqp, s_q, s_bnd,
                                    a.hash_num, L, lane);
      else
        rank_tile<W32, EUCLID, PV>(ca, (uint32_t)(t * 32 + lane), fa, nqp, s_q, s_bnd, a.hash_num, L, lane);
      if (stash_count(L, warp) >= STASH_FLUSH_AT) stash_flush<EUCLID>(warp, lane);
      if (warp == 0 && ((++poll_tick) & 7u) == 0u) poll_shared_bounds<EUCLID>(L, nqp, lane);
    }
  } else if constexpr (!REG_PREFETCH) {
    uint32_t cur[W32];
    float ncur = 0.0f;
    for (; t < t_end; t += t_step) {
      load_row<W32>(a.slab, t, lane, cur);
      if (EUCLID) ncur = ld_stream_f32(a.norms + t * 32 + lane);
      if (t + t_step < t_end) prefetch_row_l2<W32>(a.slab, t + t_step, lane);
      rank_tile<W32, EUCLID, PV>(cur, (uint32_t)(t * 32 + lane), ncur, nqp, s_q, s_bnd, a.hash_num, L, lane);
      if (stash_count(L, warp) >= STASH_FLUSH_AT) stash_flush<EUCLID>(warp, lane);
      if (warp == 0 && ((++poll_tick) & 7u) == 0u) poll_shared_bounds<EUCLID>(L, nqp, lane);
    }
  } else {
  uint32_t cur[W32], nxt[W32];
  float ncur = 0.0f, nnxt = 0.0f;
  if (t < t_end) {
    load_row<W32>(a.slab, t, lane, cur);
    if (EUCLID) ncur = ld_stream_f32(a.norms + t * 32 + lane);
  }
  for (; t < t_end; t += t_step) {
    const uint64_t tn = t + t_step;
    if (tn < t_end) {
      load_row<W32>(a.slab, tn, lane, nxt);
      if (EUCLID) nnxt = ld_stream_f32(a.norms + tn * 32 + lane);
    }
    rank_tile<W32, EUCLID, PV>(cur, (uint32_t)(t * 32 + lane), ncur, nqp, s_q, s_bnd, a.hash_num, L, lane);
    if (stash_count(L, warp) >= STASH_FLUSH_AT) stash_flush<EUCLID>(warp, lane);   // the next tile's loads are already in flight
    if (warp == 0 && ((++poll_tick) & 7u) == 0u) poll_shared_bounds<EUCLID>(L, nqp, lane);
#pragma unroll
    for (int w = 0; w < W32; ++w) cur[w] = nxt[w];
    ncur = nnxt;
  }
  }
  if (stash_count(L, warp)) stash_flush<EUCLID>(warp, lane, false);
  __syncthreads();
  for (uint32_t q = warp; q < nqp; q += SCAN_WARPS) {
    // one last look at what the whole GPU has counted: the CTA's own k-th best is a loose bound when
    // hundreds of CTAs each keep k, and everything above the common bound leaves as padding below
    if (L.ghist) hist_refresh<EUCLID>(q, lane);
    uint32_t n = list_compact<EUCLID>(q, sm_u32(L.cnt_off)[q], lane, true);
    const unsigned long long* src = sm_u64(L.list_off) + (size_t)q * a.cap;
    unsigned long long* out = reinterpret_cast<unsigned long long*>(a.cand) +
                              ((size_t)(q0 + q) * a.nchunks + blockIdx.x) * a.k;
    const unsigned long long tau = sm_u64(L.tau_off)[q];   // (what lies above the final threshold leaves as padding)
    for (uint32_t j = lane; j < a.k; j += 32) out[j] = (j < n && src[j] <= tau) ? src[j] : KEY_SENTINEL;
  }
}

// ---------------------------------------------------------------------------
// variant 2: rows staged through shared memory by the TMA engine.  One producer thread
// issues 1-D bulk copies (cp.async.bulk -> UBLKCP) of whole runs of tiles into a ring of
// stages, completion is tracked by mbarrier transaction counts; 16 consumer warps read their
// tile from shared memory with conflict-free 128-bit loads (the slab's tile-transposed
// layout is kept as is) and release the stage through a second mbarrier.  No row data
// passes through registers before it is needed, so the kernel is spill-free at 1 CTA / SM
// and the memory pipeline runs ahead of the popcount work by several stages.
// ---------------------------------------------------------------------------
#ifndef JBN_TMA_CW
#define JBN_TMA_CW 15   // 15 consumer warps + 1 producer warp = 512 threads: 128 registers per thread at 1 CTA/SM
#endif
static const int TMA_CONSUMER_WARPS = JBN_TMA_CW;
static const int TMA_THREADS = (TMA_CONSUMER_WARPS + 1) * 32;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

struct TmaPlan {
  uint32_t stages, tiles_per_warp;   // a stage holds TMA_CONSUMER_WARPS * tiles_per_warp tiles
  size_t ring_off, norm_off, bar_off, total;
};
__host__ __device__ inline TmaPlan plan_tma(uint32_t qp, uint32_t w32, uint32_t cap, bool euclid, uint32_t hash_num,
                                            uint32_t stages, uint32_t tiles_per_warp) {
  TmaPlan t;
  t.stages = stages; t.tiles_per_warp = tiles_per_warp;
  const size_t stage_tiles = (size_t)TMA_CONSUMER_WARPS * tiles_per_warp;
  size_t o = (plan_smem(qp, w32, cap, euclid, hash_num, 0, true).total + 127) & ~(size_t)127;
  t.ring_off = o; o += stages * stage_tiles * w32 * 128;
  t.norm_off = o; if (euclid) o += stages * stage_tiles * 128;
  t.bar_off = o; o += (size_t)stages * 16;
  t.total = o;
  return t;
}

template <int W32>
__device__ __forceinline__ void load_row_smem(const unsigned char* stage, uint32_t tile_in_stage, int lane,
                                              uint32_t (&r)[W32]) {
  if (W32 == 2) {
    uint2 v = *reinterpret_cast<const uint2*>(stage + ((size_t)tile_in_stage * 32 + lane) * 8);
    r[0] = v.x; r[1] = v.y;
  } else {
    constexpr int C = W32 / 4;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      uint4 v = *reinterpret_cast<const uint4*>(stage + (((size_t)tile_in_stage * C + c) * 32 + lane) * 16);
      r[4 * c + 0] = v.x; r[4 * c + 1] = v.y; r[4 * c + 2] = v.z; r[4 * c + 3] = v.w;
    }
  }
}

// named barrier of the consumer warps only (the producer warp never joins: it keeps the ring full
// while the consumers change over from one query group to the next)
__device__ __forceinline__ void consumer_barrier() {
  asm volatile("bar.sync 1, %0;" ::"r"(TMA_CONSUMER_WARPS * 32) : "memory");
}

// The CTA's candidate lists of one query group -> scratch, k per (query, CTA).  A list of up to 512
// entries is cut to its k smallest by counting: entry i's rank is the number of smaller entries (keys
// are distinct), the k smallest go to out[rank] -- no search for the k-th key, no packing, and the
// list leaves sorted.  The queries of the group are shared out over teams of warps.  Longer lists (k
// above 256) keep the one-warp selection of list_compact.
template <bool EUCLID>
__device__ __forceinline__ void write_lists(const ScanArgs& a, const ListCtl& L, uint32_t q0, uint32_t nqp, int warp, int lane,
                                            int nwarps) {
  if (a.cap <= 512) {
    const uint32_t nteams = min(nqp, (uint32_t)nwarps), wpt = (uint32_t)nwarps / nteams;
    const uint32_t team = (uint32_t)warp / wpt;
    if (team >= nteams) return;
    const uint32_t tid = ((uint32_t)warp % wpt) * 32 + lane, tsize = wpt * 32;
    for (uint32_t q = team; q < nqp; q += nteams) {
      const uint32_t cnt = sm_u32(L.cnt_off)[q];
      const unsigned long long* src = sm_u64(L.list_off) + (size_t)q * a.cap;
      unsigned long long* out = reinterpret_cast<unsigned long long*>(a.cand) + ((size_t)(q0 + q) * a.nchunks + blockIdx.x) * a.k;
      // What is above the CTA's final threshold (its own k-th best or the bound the whole GPU agreed on)
      // cannot be in the answer: it leaves as padding, and the list reduction only ever looks at the
      // few hundred keys per query that are left (merge_sparse_kernel).
      const unsigned long long tau = sm_u64(L.tau_off)[q];
      if (cnt <= a.k) {
        for (uint32_t j = tid; j < a.k; j += tsize) out[j] = (j < cnt && src[j] <= tau) ? src[j] : KEY_SENTINEL;
        continue;
      }
      for (uint32_t i = tid; i < cnt; i += tsize) {
        const unsigned long long v = src[i];
        uint32_t rank = 0;
        uint32_t j = 0;
        for (; j + 4 <= cnt; j += 4)
          rank += (src[j] < v ? 1u : 0u) + (src[j + 1] < v ? 1u : 0u) + (src[j + 2] < v ? 1u : 0u) + (src[j + 3] < v ? 1u : 0u);
        for (; j < cnt; ++j) rank += src[j] < v ? 1u : 0u;
        if (rank < a.k) out[rank] = (v <= tau) ? v : KEY_SENTINEL;   // cnt > k: every rank below k is written by somebody
      }
    }
    return;
  }
  for (uint32_t q = warp; q < nqp; q += nwarps) {
    uint32_t n = list_compact<EUCLID>(q, sm_u32(L.cnt_off)[q], lane, true);
    const unsigned long long* src = sm_u64(L.list_off) + (size_t)q * a.cap;
    unsigned long long* out = reinterpret_cast<unsigned long long*>(a.cand) + ((size_t)(q0 + q) * a.nchunks + blockIdx.x) * a.k;
    const unsigned long long tau = sm_u64(L.tau_off)[q];
    for (uint32_t j = lane; j < a.k; j += 32) out[j] = (j < n && src[j] <= tau) ? src[j] : KEY_SENTINEL;
  }
}

// Persistent over the query groups of the launch: grid = one CTA per SM; a CTA walks its share of the
// table once per group of qp queries.  The producer's ring does not stop at a group boundary -- while
// the consumers write out the lists of group g and set up the queries, thresholds and (euclid) norm
// bounds of group g + 1, the stages of g + 1 are already landing -- so a launch pays one ring fill and
// one drain, not one per group, which is what a 12.5 M-row shard (0.1 ms per pass) spent a third of
// its time on.  The share of a CTA rotates with the group, so that the CTAs that get the odd extra
// stage are not the same every time.
template <int W32, bool EUCLID, int PV, int NQ>
__global__ void __launch_bounds__(TMA_THREADS, 1) scan_tma_kernel(ScanArgs a, uint32_t stages, uint32_t tiles_per_warp) {
  unsigned char* smem = g_smem;
  const SmemPlan sp = plan_smem(a.qp, W32, a.cap, EUCLID, a.hash_num, 0, true);
  const TmaPlan tp = plan_tma(a.qp, W32, a.cap, EUCLID, a.hash_num, stages, tiles_per_warp);
  unsigned char* ring = smem + tp.ring_off;
  float* nring = reinterpret_cast<float*>(smem + tp.norm_off);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + tp.bar_off);
  uint64_t* empty_bar = full_bar + stages;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t ngroups = (a.nq + a.qp - 1) / a.qp;
  const uint32_t stage_tiles = TMA_CONSUMER_WARPS * tiles_per_warp;
  const size_t tile_bytes = (size_t)W32 * 128;
  const size_t stage_bytes = stage_tiles * tile_bytes;
  if (tid == 0) {
    for (uint32_t s = 0; s < stages; ++s) {
      mbar_init(full_bar + s, 1);
      mbar_init(empty_bar + s, TMA_CONSUMER_WARPS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  // the CTA's share of the table for group g: stage `it` starts at tile t_first + it * t_stride
  const uint64_t nstage = (a.ntiles + stage_tiles - 1) / stage_tiles;
  auto share = [&](uint32_t g, uint64_t& t_first, uint64_t& t_stride, uint64_t& t_end, uint32_t& iters) {
    const uint32_t c = (blockIdx.x + g) % gridDim.x;
    if (a.interleave) {
      t_first = (uint64_t)c * stage_tiles;
      t_stride = (uint64_t)gridDim.x * stage_tiles;
      t_end = a.ntiles;
      iters = nstage > c ? (uint32_t)((nstage - c + gridDim.x - 1) / gridDim.x) : 0u;
    } else {
      t_first = (uint64_t)c * a.tiles_per_chunk;
      t_stride = stage_tiles;
      t_end = min(a.ntiles, t_first + (uint64_t)a.tiles_per_chunk);
      const uint64_t my_tiles = t_end > t_first ? t_end - t_first : 0;
      iters = (uint32_t)((my_tiles + stage_tiles - 1) / stage_tiles);
    }
  };

  if (warp == TMA_CONSUMER_WARPS) {
    // ---- producer: one elected thread feeds the ring, group after group without a pause ----
    if (lane == 0) {
      uint32_t s = 0, ph = 0, issued = 0;
      for (uint32_t g = 0; g < ngroups; ++g) {
        uint64_t t0, t_stride, t_end;
        uint32_t iters;
        share(g, t0, t_stride, t_end, iters);
        for (uint32_t it = 0; it < iters; ++it, t0 += t_stride, ++issued) {
          if (issued >= stages) mbar_wait(empty_bar + s, ph ^ 1u);
          const uint32_t ntile = (uint32_t)min((uint64_t)stage_tiles, t_end - t0);
          const uint32_t bytes = ntile * (uint32_t)tile_bytes;
          mbar_expect_tx(full_bar + s, bytes + (EUCLID ? ntile * 128u : 0u));
          bulk_g2s(ring + s * stage_bytes, a.slab + t0 * tile_bytes, bytes, full_bar + s);
          if (EUCLID) bulk_g2s(nring + (size_t)s * stage_tiles * 32, a.norms + t0 * 32, ntile * 128u, full_bar + s);
          if (++s == stages) { s = 0; ph ^= 1u; }
        }
      }
    }
    __syncwarp();
    return;
  }

  // ---- consumers: warp w owns tiles 2w, 2w+1 (+ 2*15 per further pair) of every stage ----
  constexpr int CT = TMA_CONSUMER_WARPS * 32;
  constexpr int NQR = NQ > 0 ? NQ : 1;
  constexpr uint32_t QS = q_stride(W32);
  uint32_t s = 0, ph = 0;
  ListCtl L;
  uint32_t* s_q;
  float* s_bnd;
  if (EUCLID)   // the cos table serves every group (visible after the first group's prologue barrier)
    for (uint32_t d = tid; d <= a.hash_num; d += CT) reinterpret_cast<double*>(smem + sp.cos_off)[d] = a.coslut[d];
  for (uint32_t g = 0; g < ngroups; ++g) {
    const uint32_t q0 = g * a.qp;
    const uint32_t nqp = min(a.qp, a.nq - q0);
#ifdef JBN_SCAN_TRACE   // tuning builds only (tools/scan_trace.py): the stage loop pays for every test in it
#define JBN_TRACE(slot)                                                                                     \
  if (a.trace && blockIdx.x < 4 && tid == 0) {                                                              \
    unsigned long long t_;                                                                                  \
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                                  \
    a.trace[((size_t)blockIdx.x * ngroups + g) * 8 + (slot)] = t_;                                          \
  }
#else
#define JBN_TRACE(slot)
#endif
    JBN_TRACE(0)
    scan_prologue<W32, EUCLID, true>(a, smem, sp, L, s_q, s_bnd, q0, nqp, tid, CT);
    consumer_barrier();
    JBN_TRACE(1)
    uint64_t t0, t_stride, t_end;
    uint32_t iters;
    share(g, t0, t_stride, t_end, iters);
    uint32_t qreg[NQR][W32];
    if (NQ > 0) {
#pragma unroll
      for (int q = 0; q < NQR; ++q) load_query<W32>(s_q + q * QS, qreg[q]);
    }
    uint32_t next_refresh = (L.ghist && !(a.debug & 1u)) ? 0u : 0xFFFFFFFFu;
    const uint32_t poll_mask = (L.ghist && !(a.debug & 4u)) ? (EUCLID ? 7u : 3u) : 0xFFFFFFFFu;   // (all ones: never equal below)
    uint32_t poll_warp = 0;
    for (uint32_t it = 0; it < iters; ++it, t0 += t_stride) {
      mbar_wait(full_bar + s, ph);
      const unsigned char* stage = ring + s * stage_bytes;
      const float* nstage_p = nring + (size_t)s * stage_tiles * 32;
      if (NQ > 0) {
        uint32_t tau[NQR];
#pragma unroll
        for (int q = 0; q < NQR; ++q) tau[q] = *reinterpret_cast<const volatile uint32_t*>(s_q + q * QS + W32);
        for (uint32_t j = 0; j < tiles_per_warp; j += 2) {
          const uint32_t tis = (j * TMA_CONSUMER_WARPS) + 2 * warp;
          const uint64_t t = t0 + tis;
          if (t >= t_end) break;
          const uint32_t rowa = (uint32_t)(t * 32 + lane);
          if (NQ <= 2 && t + 1 < t_end) {
            uint32_t rows[2][W32];
            load_row_smem<W32>(stage, tis, lane, rows[0]);
            load_row_smem<W32>(stage, tis + 1, lane, rows[1]);
            const uint32_t rowid[2] = {rowa, rowa + 32};
            float ni[2] = {0.0f, 0.0f};
            if (EUCLID) { ni[0] = nstage_p[tis * 32 + lane]; ni[1] = nstage_p[(tis + 1) * 32 + lane]; }
            rank_reg<W32, EUCLID, PV, NQR, 2>(rows, rowid, ni, qreg, tau, s_bnd, a.hash_num, warp, lane);
          } else {
            const uint32_t ntile = (t + 1 < t_end) ? 2u : 1u;
            for (uint32_t u = 0; u < ntile; ++u) {
              uint32_t rows[1][W32];
              load_row_smem<W32>(stage, tis + u, lane, rows[0]);
              const uint32_t rowid[1] = {rowa + 32 * u};
              float ni[1] = {0.0f};
              if (EUCLID) ni[0] = nstage_p[(tis + u) * 32 + lane];
              rank_reg<W32, EUCLID, PV, NQR, 1>(rows, rowid, ni, qreg, tau, s_bnd, a.hash_num, warp, lane);
            }
          }
        }
      } else
      for (uint32_t j = 0; j < tiles_per_warp; j += 2) {
        const uint32_t tis = (j * TMA_CONSUMER_WARPS) + 2 * warp;
        const uint64_t t = t0 + tis;
        if (t >= t_end) break;
        uint32_t ra[W32];
        load_row_smem<W32>(stage, tis, lane, ra);
        float nia = 0.0f;
        if (EUCLID) nia = nstage_p[tis * 32 + lane];
        const uint32_t rowa = (uint32_t)(t * 32 + lane);
        if (t + 1 < t_end) {
          uint32_t rb[W32];
          load_row_smem<W32>(stage, tis + 1, lane, rb);
          float nib = 0.0f;
          if (EUCLID) nib = nstage_p[(tis + 1) * 32 + lane];
          rank_tile2<W32, EUCLID, PV>(ra, rb, rowa, rowa + 32, nia, nib, nqp, s_q, s_bnd, a.hash_num, L, lane);
        } else {
          rank_tile<W32, EUCLID, PV>(ra, rowa, nia, nqp, s_q, s_bnd, a.hash_num, L, lane);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty_bar + s);   // release the stage first ...
      if (stash_count(L, warp) >= STASH_FLUSH_AT) stash_flush<EUCLID>(warp, lane);   // ... then do the slow insertions
      // euclid: a tighter bound means recomputing that query's norm bounds, so poll less often
      // and take turns (a warp that lags holds up the whole ring)
      // bounds maintenance, kept to one compare per stage on the common path (46 more instructions per
      // warp and stage here cost the 10^8-row headline 7 %)
      if (it == next_refresh) {
        // start of a pass: the thresholds come from a small sample and are loose; after one stage per
        // CTA the histogram already holds far more than k candidates, so one warp per stage (in turn)
        // derives the bound from it directly instead of waiting for somebody's flush to publish one.
        // Later candidates are so rare that nobody flushes any more: the bound is re-derived at stages
        // 4, 8, 16, ... (hamming engines: one short read; a euclid_lsh refresh costs a warp two
        // microseconds, which a pass that is exactly as fast as HBM has no room for) and once more three
        // stages before the end, so that what the lists hold above the final bound (blanked in
        // write_lists) is a few keys, not thousands.
        // (query q by warp (stage + q) mod 15: no warp is held up by more than one refresh)
        const uint32_t slot = ((uint32_t)warp + TMA_CONSUMER_WARPS - it % TMA_CONSUMER_WARPS) % TMA_CONSUMER_WARPS;
        for (uint32_t q = slot; q < nqp; q += TMA_CONSUMER_WARPS) hist_refresh<EUCLID>(q, lane);
        uint32_t nx = 0xFFFFFFFFu;
        if (it + 1 < EARLY_REFRESH_STAGES) nx = it + 1;
        else if (!EUCLID) nx = 1u << (32 - __clz(it));   // the next power of two above it
        if (iters >= 3 && iters - 3 > it && iters - 3 < nx) nx = iters - 3;
        next_refresh = nx;
      } else if ((it & poll_mask) == poll_mask) {
        // published bounds (one word per query from L2): every 4th stage for the hamming engines; for
        // euclid_lsh, where a tighter bound means recomputing that query's norm bounds, every 8th, the
        // warps taking turns (a warp that lags holds up the whole ring)
        if ((uint32_t)warp == poll_warp) poll_shared_bounds<EUCLID>(L, nqp, lane);
        if (EUCLID) poll_warp = poll_warp + 1 == TMA_CONSUMER_WARPS ? 0u : poll_warp + 1;   // (every warp keeps the same count)
      }
      if (++s == stages) { s = 0; ph ^= 1u; }
#ifdef JBN_SCAN_TRACE
      if (it == 0) { JBN_TRACE(2) } else if (it == 3) { JBN_TRACE(3) } else if (it == 11) { JBN_TRACE(4) }
#endif
    }
    JBN_TRACE(5)
    if (stash_count(L, warp)) stash_flush<EUCLID>(warp, lane, false);   // the pass is over: a tighter bound serves nobody
    consumer_barrier();
    JBN_TRACE(6)
    write_lists<EUCLID>(a, L, q0, nqp, warp, lane, TMA_CONSUMER_WARPS);
    consumer_barrier();   // the lists are out: the next group's prologue may overwrite them
    JBN_TRACE(7)
  }
}

// ---------------------------------------------------------------------------
// merge of the per-CTA lists of one query: k smallest of nlists*k keys
// ---------------------------------------------------------------------------
__device__ __forceinline__ void block_bitonic_sort(unsigned long long* a, uint32_t n, int tid, int nthreads) {
  for (uint32_t size = 2; size <= n; size <<= 1) {
    for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
      for (uint32_t i = tid; i < (n >> 1); i += nthreads) {
        uint32_t lo = ((i & ~(stride - 1)) << 1) | (i & (stride - 1));
        uint32_t hi = lo | stride;
        bool asc = ((lo & size) == 0);
        unsigned long long x = a[lo], y = a[hi];
        if ((x > y) == asc && x != y) { a[lo] = y; a[hi] = x; }
      }
      __syncthreads();
    }
  }
}

__device__ __forceinline__ float final_score(uint32_t key_hi, const MergeOut& o, uint32_t q) {
  float s;
  if (o.euclid) {
    // sqrt(norm_q^2 + score): float mul, float add, correctly rounded sqrt
    const float nq = o.qnorms[q];
    s = __fsqrt_rn(__fadd_rn(__fmul_rn(nq, nq), key_float(key_hi)));
    if (o.flags & 1u) s = -s;
  } else {
    s = __fdiv_rn((float)key_hi, (float)o.hash_num);
    if (o.flags & 1u) s = __fsub_rn(1.0f, s);
  }
  return s;
}

// half: elements kept between sorts (>= k, >= MERGE_THREADS*2, power of two); smem holds 2*half keys
__global__ void __launch_bounds__(MERGE_THREADS) merge_local_kernel(const unsigned long long* cand,
                                                                    uint32_t nlists, uint32_t k,
                                                                    uint32_t half, MergeOut o, const uint32_t* only_dense) {
  extern __shared__ __align__(16) unsigned char smem[];
  if (only_dense && !only_dense[blockIdx.x]) return;   // merge_sparse_kernel has answered this query
  unsigned long long* buf = reinterpret_cast<unsigned long long*>(smem);
  __shared__ uint32_t s_n;
  const int tid = threadIdx.x;
  const uint32_t q = blockIdx.x;
  const unsigned long long* src = cand + (size_t)q * nlists * k;
  const size_t total = (size_t)nlists * k;
  for (uint32_t i = tid; i < 2 * half; i += MERGE_THREADS) buf[i] = KEY_SENTINEL;
  if (tid == 0) s_n = 0;
  __syncthreads();
  unsigned long long tau = KEY_SENTINEL;  // k-th smallest key so far (uniform across the CTA)
  uint32_t pending = 0;                   // keys waiting in buf[half..), uniform
  for (size_t base = 0; base < total; base += MERGE_THREADS) {
    const size_t i = base + tid;
    bool take = false;
    if (i < total) {
      unsigned long long key = src[i];
      take = key < tau;
      if (take) {
        uint32_t pos = atomicAdd(&s_n, 1u);
        buf[half + pos] = key;
      }
    }
    pending += __syncthreads_count(take);
    const bool last = (base + MERGE_THREADS >= total);
    if (pending + MERGE_THREADS > half || last) {
      for (uint32_t j = half + pending + tid; j < 2 * half; j += MERGE_THREADS) buf[j] = KEY_SENTINEL;
      __syncthreads();
      block_bitonic_sort(buf, 2 * half, tid, MERGE_THREADS);
      tau = buf[k - 1];
      pending = 0;
      if (tid == 0) s_n = 0;
      __syncthreads();
    }
  }
  // buf[0..k) ascending
  uint32_t count = 0;
  for (uint32_t j = tid; j < k; j += MERGE_THREADS) {
    const unsigned long long key = buf[j];
    const bool ok = key != KEY_SENTINEL;
    const uint32_t key_hi = (uint32_t)(key >> 32);
    const uint64_t gidx = ok ? (uint64_t)(uint32_t)key * o.shard_count + o.shard_rank : IDX_SENTINEL;
    if (o.cand_key && ok) {  // unused slots were pre-filled with sentinels by the host
      o.cand_key[(size_t)q * o.kout + o.col0 + j] = key_hi;
      o.cand_index[(size_t)q * o.kout + o.col0 + j] = gidx;
    }
    if (o.out_index && ok) {
      o.out_index[(size_t)q * o.kout + o.col0 + j] = gidx;
      o.out_score[(size_t)q * o.kout + o.col0 + j] = final_score(key_hi, o, q);
    }
    count += ok ? 1u : 0u;
  }
  // count valid entries
  __shared__ uint32_t s_count;
  if (tid == 0) s_count = 0;
  __syncthreads();
  if (count) atomicAdd(&s_count, count);
  __syncthreads();
  if (tid == 0) {
    if (o.out_count) o.out_count[q] = (o.count_accumulate ? o.out_count[q] : 0u) + s_count;
    if (o.last_key) o.last_key[q] = s_count ? buf[s_count - 1] : KEY_SENTINEL;
  }
}

// ---------------------------------------------------------------------------
// merge of `parts` shard-local candidate lists ([parts][nq][k], each ascending by
// (key, global index), sentinel padded): rank of an element = its position in its own list
// + number of smaller elements in every other list (binary search); rank < k -> output.
// ---------------------------------------------------------------------------
__device__ __forceinline__ bool cand_less(uint32_t ka, uint64_t ia, uint32_t kb, uint64_t ib) {
  return ka < kb || (ka == kb && ia < ib);
}

// part p's lists start at ckey + p * key_stride and cidx + p * idx_stride (elements): separate
// [parts][nq][k] arrays, or the slots of one all-gathered buffer that packs both per rank
__global__ void __launch_bounds__(MERGE_THREADS) merge_parts_kernel(const uint32_t* ckey, const uint64_t* cidx,
                                                                    uint32_t nq, uint32_t k, uint32_t parts,
                                                                    size_t key_stride, size_t idx_stride, MergeOut o) {
  const uint32_t q = blockIdx.x;
  __shared__ uint32_t s_count;
  if (threadIdx.x == 0) s_count = 0;
  __syncthreads();
  uint32_t mine = 0;
  for (uint32_t e = threadIdx.x; e < parts * k; e += MERGE_THREADS) {
    const uint32_t p = e / k, j = e % k;
    const size_t at = (size_t)q * k + j;
    const uint32_t key = ckey[(size_t)p * key_stride + at];
    const uint64_t idx = cidx[(size_t)p * idx_stride + at];
    if (idx == IDX_SENTINEL) continue;
    uint32_t rank = j;
    for (uint32_t p2 = 0; p2 < parts && rank < k; ++p2) {
      if (p2 == p) continue;
      const uint32_t* k2 = ckey + (size_t)p2 * key_stride + (size_t)q * k;
      const uint64_t* i2 = cidx + (size_t)p2 * idx_stride + (size_t)q * k;
      uint32_t lo = 0, hi = k;  // first element of list p2 that is not less than (key, idx)
      while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        if (cand_less(k2[mid], i2[mid], key, idx)) lo = mid + 1; else hi = mid;
      }
      rank += lo;
    }
    if (rank < k) {
      o.out_index[(size_t)q * o.kout + o.col0 + rank] = idx;
      o.out_score[(size_t)q * o.kout + o.col0 + rank] = final_score(key, o, q);
      ++mine;
    }
  }
  if (mine) atomicAdd(&s_count, mine);
  __syncthreads();
  if (threadIdx.x == 0 && o.out_count) o.out_count[q] = s_count;
}

// List reduction when all nlists * k candidates of a query fit in registers (up to 24576 keys):
// the k-th smallest key by a most-significant-bit-first search -- a round is KPT compares, one
// warp reduction, one shared atomic and one barrier (three counters in rotation), and the bits
// on which all keys agree (unused high bits of the row index, most of the distance) cost no
// round -- then the k survivors are ordered by counting (each finds its rank among the others)
// or, for large k, by a bitonic sort.  The streaming variant above sorts 2 * half keys per 256
// candidates read, which is what a single small query spent most of its time in.
static const uint32_t MSEL_MAX_KEYS = 24576;

// List reduction for the common case: the scan CTAs have already blanked everything above their final
// thresholds, so of the nlists * k slots of a query only a few hundred hold keys.  They are collected
// in shared memory and ordered by counting (keys are distinct: the number of smaller keys is the
// position).  A query with more than MSP_CAP live keys (long runs of equal distances) is flagged in
// `dense` and left to the selection kernels below, which skip the queries that are not flagged.
static const uint32_t MSP_CAP = 1024, MSP_THREADS = 512;   // (ordering by counting is quadratic: 1024 keys = 2 x 1024 compares per thread)
__global__ void __launch_bounds__(MSP_THREADS) merge_sparse_kernel(const unsigned long long* cand, uint32_t nlists, uint32_t k,
                                                                   MergeOut o, uint32_t* dense) {
  __shared__ unsigned long long buf[MSP_CAP];
  __shared__ uint32_t s_n;
  const uint32_t q = blockIdx.x, tid = threadIdx.x;
  const size_t M = (size_t)nlists * k;
  const unsigned long long* src = cand + (size_t)q * M;
  if (tid == 0) s_n = 0;
  __syncthreads();
  for (size_t i = tid; i < M; i += MSP_THREADS) {
    const unsigned long long v = src[i];
    if (v != KEY_SENTINEL) {
      const uint32_t pos = atomicAdd(&s_n, 1u);
      if (pos < MSP_CAP) buf[pos] = v;
    }
  }
  __syncthreads();
  const uint32_t n = s_n;
  if (n > MSP_CAP) { if (tid == 0) dense[q] = 1u; return; }
  if (tid == 0) dense[q] = 0u;
  const uint32_t want = min(k, n);
  for (uint32_t i = tid; i < n; i += MSP_THREADS) {
    const unsigned long long key = buf[i];
    uint32_t rank = 0, j = 0;
    for (; j + 4 <= n; j += 4)
      rank += (buf[j] < key ? 1u : 0u) + (buf[j + 1] < key ? 1u : 0u) + (buf[j + 2] < key ? 1u : 0u) + (buf[j + 3] < key ? 1u : 0u);
    for (; j < n; ++j) rank += buf[j] < key ? 1u : 0u;
    if (rank >= want) continue;
    const uint32_t key_hi = (uint32_t)(key >> 32);
    const uint64_t gidx = (uint64_t)(uint32_t)key * o.shard_count + o.shard_rank;
    if (o.cand_key) {  // unused slots were pre-filled with sentinels by the host
      o.cand_key[(size_t)q * o.kout + o.col0 + rank] = key_hi;
      o.cand_index[(size_t)q * o.kout + o.col0 + rank] = gidx;
    }
    if (o.out_index) {
      o.out_index[(size_t)q * o.kout + o.col0 + rank] = gidx;
      o.out_score[(size_t)q * o.kout + o.col0 + rank] = final_score(key_hi, o, q);
    }
    if (rank == want - 1 && o.last_key) o.last_key[q] = key;
  }
  if (tid == 0) {
    if (o.out_count) o.out_count[q] = (o.count_accumulate ? o.out_count[q] : 0u) + want;
    if (o.last_key && want == 0) o.last_key[q] = KEY_SENTINEL;
  }
}

template <int KPT, int THREADS>
__global__ void __launch_bounds__(THREADS) merge_small_kernel(const unsigned long long* cand, uint32_t nlists,
                                                              uint32_t k, uint32_t kp, MergeOut o, const uint32_t* only_dense) {
  if (only_dense && !only_dense[blockIdx.x]) return;   // merge_sparse_kernel has answered this query
  extern __shared__ __align__(16) unsigned char smem[];
  unsigned long long* top = reinterpret_cast<unsigned long long*>(smem);   // [kp], kp = pow2 >= k
  __shared__ uint32_t s_cnt[3];
  __shared__ uint32_t s_pos, s_valid;
  __shared__ unsigned long long s_or, s_and;
  const int tid = threadIdx.x, lane = tid & 31;
  const uint32_t q = blockIdx.x;
  const uint32_t M = nlists * k;
  const unsigned long long* src = cand + (size_t)q * M;
  unsigned long long r[KPT];
#pragma unroll
  for (int j = 0; j < KPT; ++j) {
    const uint32_t i = j * THREADS + tid;
    r[j] = i < M ? src[i] : KEY_SENTINEL;
  }
  unsigned long long vor = 0, vand = ~0ull;
  uint32_t valid = 0;
#pragma unroll
  for (int j = 0; j < KPT; ++j)
    if (r[j] != KEY_SENTINEL) { vor |= r[j]; vand &= r[j]; ++valid; }
  for (uint32_t i = tid; i < kp; i += THREADS) top[i] = KEY_SENTINEL;
  if (tid == 0) { s_cnt[0] = 0; s_cnt[1] = 0; s_cnt[2] = 0; s_pos = 0; s_or = 0; s_and = ~0ull; s_valid = 0; }
#pragma unroll
  for (int ofs = 16; ofs > 0; ofs >>= 1) {
    vor |= __shfl_xor_sync(FULL_MASK, vor, ofs);
    vand &= __shfl_xor_sync(FULL_MASK, vand, ofs);
  }
  valid = __reduce_add_sync(FULL_MASK, valid);
  __syncthreads();
  if (lane == 0 && valid) { atomicOr(&s_or, vor); atomicAnd(&s_and, vand); atomicAdd(&s_valid, valid); }
  __syncthreads();
  const uint32_t nvalid = s_valid;
  const uint32_t want = min(k, nvalid);
  unsigned long long kth = KEY_SENTINEL;
  if (nvalid > k) {
    const unsigned long long all_or = s_or, all_and = s_and, varying = all_or ^ all_and;
    unsigned long long prefix = 0;
    uint32_t round = 0;
    for (int bit = 63 - __clzll((long long)(all_or | 1ull)); bit >= 0; --bit) {
      if (!((varying >> bit) & 1ull)) { prefix |= all_and & (1ull << bit); continue; }
      const unsigned long long trial = prefix | (1ull << bit);
      uint32_t c = 0;
#pragma unroll
      for (int j = 0; j < KPT; ++j) c += (r[j] < trial) ? 1u : 0u;
      c = __reduce_add_sync(FULL_MASK, c);
      const uint32_t slot = round % 3u;
      if (lane == 0 && c) atomicAdd(&s_cnt[slot], c);
      if (tid == 0) s_cnt[(round + 1u) % 3u] = 0;   // last read two rounds ago, behind a barrier
      __syncthreads();
      if (s_cnt[slot] < k) prefix = trial;
      ++round;
    }
    kth = prefix;
  }
#pragma unroll
  for (int j = 0; j < KPT; ++j)
    if (r[j] != KEY_SENTINEL && r[j] <= kth) top[atomicAdd(&s_pos, 1u)] = r[j];   // exactly `want` keys
  __syncthreads();
  unsigned long long* sorted = top;
  if (kp <= 512) {
    // order by counting: keys are distinct, so the number of smaller survivors is the position
    unsigned long long* ord = top + kp;
    for (uint32_t i = tid; i < want; i += THREADS) {
      const unsigned long long v = top[i];
      uint32_t rank = 0;
      for (uint32_t j = 0; j < want; ++j) rank += (top[j] < v) ? 1u : 0u;
      ord[rank] = v;
    }
    __syncthreads();
    sorted = ord;
  } else {
    block_bitonic_sort(top, kp, tid, THREADS);
  }
  for (uint32_t j = tid; j < want; j += THREADS) {
    const unsigned long long key = sorted[j];
    const uint32_t key_hi = (uint32_t)(key >> 32);
    const uint64_t gidx = (uint64_t)(uint32_t)key * o.shard_count + o.shard_rank;
    if (o.cand_key) {  // unused slots were pre-filled with sentinels by the host
      o.cand_key[(size_t)q * o.kout + o.col0 + j] = key_hi;
      o.cand_index[(size_t)q * o.kout + o.col0 + j] = gidx;
    }
    if (o.out_index) {
      o.out_index[(size_t)q * o.kout + o.col0 + j] = gidx;
      o.out_score[(size_t)q * o.kout + o.col0 + j] = final_score(key_hi, o, q);
    }
  }
  if (tid == 0) {
    if (o.out_count) o.out_count[q] = (o.count_accumulate ? o.out_count[q] : 0u) + want;
    if (o.last_key) o.last_key[q] = want ? sorted[want - 1] : KEY_SENTINEL;
  }
}


// ---------------------------------------------------------------------------
// sample pre-pass: an admissible-key bound per query before the scan starts.
// One CTA per query ranks the first S rows and takes the k-th smallest ordering key t;
// at least k rows have key_hi <= t, so every top-k key is below (t + 1) << 32.  Starting the
// scan with this bound removes the warm-up in which every row of a CTA's first tiles is a
// candidate and all warps queue on the list lock.  (tau_init_hamming_kernel and
// tau_init_euclid_kernel, further down next to their launch site.)
// ---------------------------------------------------------------------------

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
uint32_t max_k_per_pass() { return MAX_CAP / 2; }
uint32_t hist_stride_for(uint32_t hash_num, bool euclid) {
  return euclid ? EUCLID_COARSE_BINS + EUCLID_FINE_BINS : (hash_num + 1 + 31) / 32 * 32;
}
size_t hist_scratch_bytes(uint32_t nq, uint32_t hash_num, bool euclid) {
  return ((size_t)nq + (size_t)nq * hist_stride_for(hash_num, euclid)) * 4;
}

static uint32_t g_tma_tpw_override = 0, g_tma_stages_override = 0;
void set_tma_geometry_override(uint32_t tiles_per_warp, uint32_t stages) {
  g_tma_tpw_override = tiles_per_warp; g_tma_stages_override = stages;
}

static uint32_t list_cap_for(uint32_t k) {
  uint32_t need = k < 32 ? 64 : 2 * k;
  uint32_t cap = 64;
  while (cap < need) cap <<= 1;
  return cap;
}

bool plan_scan(const TableView& t, bool euclid, uint32_t nq, uint32_t k, int variant, int popc_variant,
               uint32_t qp_opt, int interleave, int num_sms, ScanConfig* cfg) {
  if (k == 0 || k > max_k_per_pass()) return false;
  cfg->popc_variant = popc_variant < 0 ? 0 : popc_variant;
  cfg->cap = list_cap_for(k);
  const uint64_t ntiles = (t.local_rows + 31) / 32;
  // queries sharing one pass: by default as many as amortise the HBM stream (the scan is then
  // bound by the popcount pipes); callers that want the HBM-bound regime ask for 1..2
  // measured best on 10^8-row tables (euclid keeps a bound table per query).  Rows of 512 bits and more run
  // one CTA per SM with twice the shared memory, and 64 queries per pass beat 32 at every table size
  // (4467 against 4362 queries/s on 10^8 rows, 277 k against 255 k on a 1.25 M-row shard: a pass's fixed
  // costs are shared by twice the queries)
  // (rows of up to 128 bits likewise, where the lists fit: minhash 128 bits, k=50: 63.2 k -> 66.0 k queries/s on
  // 2*10^7 rows, 272 k -> 318 k on a 2.5 M-row shard; 256-bit rows stay at 32, euclid_lsh at 24)
  uint32_t qp = qp_opt ? qp_opt : (euclid ? 24 : ((t.w32 >= 16 || t.w32 <= 4) ? 64 : 32));
  if (qp > nq) qp = nq;
  if (qp < 1) qp = 1;
  // automatic variant: the TMA ring wins while the pass is memory-bound, the LDG kernel's
  // leaner inner loop wins once it is popcount-bound (measured crossover at 3-4 queries per pass)
  // (a table that cannot give every SM a stage or two is not worth the ring's set-up either)
  const uint32_t tma_max_qp = t.w32 >= 16 ? 3 : 4;   // 512-bit rows: the LDG kernel is ahead from 4 queries per pass on
  cfg->variant = variant <= 0 ? ((qp <= tma_max_qp && ntiles >= (uint64_t)num_sms * 32) ? 2 : 1) : variant;
  // tables of up to ~10^6 short rows: a CTA of the LDG kernel gets a handful of tiles per warp and is at
  // its fixed costs, the persistent ring does better whatever the number of queries (31 stored rows
  // re-scored on a 10^6 x 256-bit table: 152 -> 132 us per call) -- provided all of them fit one group
  if (variant <= 0 && cfg->variant == 1 && t.w32 <= 8 && ntiles >= (uint64_t)num_sms * 32 && ntiles < (uint64_t)num_sms * 256) {
    uint32_t tpw = 2;
    while ((size_t)TMA_CONSUMER_WARPS * tpw * t.w32 * 128 < 48 * 1024) tpw <<= 1;
    if (plan_tma(qp, t.w32, cfg->cap, euclid, t.hash_num, 2, tpw).total <= (size_t)226 * 1024) cfg->variant = 2;
  }
  cfg->stages = 0; cfg->tiles_per_warp = 0; cfg->score_mode = 0;
  cfg->interleave = interleave ? 1u : 0u;
  uint64_t max_chunks;
  if (cfg->variant == 2) {
    // one CTA per SM; ring of >= 2 stages of ~32 KB; the rest of shared memory for the lists
    const size_t budget = 226 * 1024;
    uint32_t tpw = 2;  // consumer warps take tiles in pairs
    while ((size_t)TMA_CONSUMER_WARPS * tpw * t.w32 * 128 < 48 * 1024) tpw <<= 1;
    while (tpw > 2 && plan_tma(qp, t.w32, cfg->cap, euclid, t.hash_num, 2, tpw).total > budget) tpw >>= 1;
    while (qp > 1 && plan_tma(qp, t.w32, cfg->cap, euclid, t.hash_num, 2, tpw).total > budget) --qp;
    if (plan_tma(qp, t.w32, cfg->cap, euclid, t.hash_num, 2, tpw).total > budget) {
      // the ring does not fit beside the lists (very wide rows / very large k): stream with LDG instead
      return plan_scan(t, euclid, nq, k, 1, popc_variant, qp_opt, interleave, num_sms, cfg);
    }
    if (g_tma_tpw_override >= 2 && plan_tma(qp, t.w32, cfg->cap, euclid, t.hash_num, 2, g_tma_tpw_override & ~1u).total <= budget)
      tpw = g_tma_tpw_override & ~1u;
    uint32_t stages = 2;
    const uint32_t max_stages = g_tma_stages_override >= 2 ? g_tma_stages_override : 6;
    while (stages < max_stages && plan_tma(qp, t.w32, cfg->cap, euclid, t.hash_num, stages + 1, tpw).total <= budget) ++stages;
    cfg->stages = stages; cfg->tiles_per_warp = tpw;
    cfg->smem_bytes = plan_tma(qp, t.w32, cfg->cap, euclid, t.hash_num, stages, tpw).total;
    max_chunks = (uint64_t)num_sms;
  } else {
    // two CTAs per SM => ~100 KB of shared memory each (rows of 512 bits and more: one CTA, twice that)
    const size_t budget = (t.w32 >= 16 ? 200 : 100) * 1024;
    while (qp > 1 && plan_smem(qp, t.w32, cfg->cap, euclid, t.hash_num, 0).total > budget) --qp;
    cfg->smem_bytes = plan_smem(qp, t.w32, cfg->cap, euclid, t.hash_num, 0).total;
    // rows of 512 bits and more run one CTA per SM: one chunk per SM, so that a query group is one wave
    // of CTAs (two waves of half-size chunks cost a 12.5 M-row shard 4.6 % at 32 queries per pass:
    // 30.9 k against 32.3 k queries/s; no difference at 10^8 rows)
    max_chunks = (uint64_t)num_sms * (t.w32 >= 16 ? 1 : 2);
  }
  cfg->qp = qp;
  // at least 64 tiles per CTA; tables too small to give every SM that much are cut finer (down
  // to one tile per warp) so that a single query is not left to a handful of SMs
  uint64_t min_tiles = 64;
  if (ntiles < (uint64_t)num_sms * 64) min_tiles = cfg->variant == 2 ? 32 : 16;
  uint64_t nchunks = (ntiles + min_tiles - 1) / min_tiles;
  if (nchunks > max_chunks) nchunks = max_chunks;
  if (nchunks < 1) nchunks = 1;
  cfg->nchunks = (uint32_t)nchunks;
  cfg->tiles_per_chunk = (uint32_t)((ntiles + nchunks - 1) / nchunks);
  if (cfg->tiles_per_chunk == 0) cfg->tiles_per_chunk = 1;
  return true;
}

template <int W32, bool EUCLID, int PV, int NQ>
static cudaError_t launch_tma_t(const ScanArgs& a, const ScanConfig& cfg, dim3 grid, cudaStream_t s) {
  auto kern = scan_tma_kernel<W32, EUCLID, PV, NQ>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem_bytes);
  if (e != cudaSuccess) return e;
  grid.y = 1;   // persistent over the query groups
  kern<<<grid, TMA_THREADS, cfg.smem_bytes, s>>>(a, cfg.stages, cfg.tiles_per_warp);
  return cudaGetLastError();
}

template <int W32, bool EUCLID, int PV>
static cudaError_t launch_scan_t(const ScanArgs& a, const ScanConfig& cfg, dim3 grid, cudaStream_t s) {
  if (cfg.variant == 2) {
    // register-resident queries for the common signature widths when every group is full
    if ((W32 == 8 || W32 == 16) && (PV == 1 || PV == 3) && a.qp <= 4 && a.nq % a.qp == 0) {
      constexpr int W = (W32 == 8 || W32 == 16) ? W32 : 16;
      constexpr int P = (PV == 1 || PV == 3) ? PV : 1;
      switch (a.qp) {
        case 1: return launch_tma_t<W, EUCLID, P, 1>(a, cfg, grid, s);
        case 2: return launch_tma_t<W, EUCLID, P, 2>(a, cfg, grid, s);
        case 3: return launch_tma_t<W, EUCLID, P, 3>(a, cfg, grid, s);
        default: return launch_tma_t<W, EUCLID, P, 4>(a, cfg, grid, s);
      }
    }
    return launch_tma_t<W32, EUCLID, PV, 0>(a, cfg, grid, s);
  }
  auto kern = scan_ldg_kernel<W32, EUCLID, PV>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem_bytes);
  if (e != cudaSuccess) return e;
  kern<<<grid, SCAN_THREADS, cfg.smem_bytes, s>>>(a);
  return cudaGetLastError();
}

template <int W32, bool EUCLID>
static cudaError_t launch_scan_pv(const ScanArgs& a, const ScanConfig& cfg, dim3 grid, cudaStream_t s) {
  switch (cfg.popc_variant) {
    case 0: return launch_scan_t<W32, EUCLID, 0>(a, cfg, grid, s);
    case 1: return launch_scan_t<W32, EUCLID, 1>(a, cfg, grid, s);
    case 2: return launch_scan_t<W32, EUCLID, 2>(a, cfg, grid, s);
    case 3: return launch_scan_t<W32, EUCLID, 3>(a, cfg, grid, s);
    default: return launch_scan_t<W32, EUCLID, 4>(a, cfg, grid, s);
  }
}

// Hamming engines: the k-th smallest distance of the sample needs no sort, distances are small
// integers -- count them in shared memory and walk the counts.  The same CTA also resets its
// query's part of the shared histogram (saves two memset nodes per launch on the latency path).
static const int TAUH_THREADS = 1024;
__global__ void __launch_bounds__(TAUH_THREADS) tau_init_hamming_kernel(ScanArgs a, uint32_t S, unsigned long long* tau_out) {
  __shared__ uint32_t hist[1025 + 31];
  __shared__ uint32_t s_found;
  const uint32_t q = blockIdx.x;
  const uint32_t bins = a.hash_num + 1;
  for (uint32_t i = threadIdx.x; i < bins + 31; i += TAUH_THREADS) hist[i] = 0;
  if (threadIdx.x == 0) s_found = 0xFFFFFFFFu;
  if (a.ghist) {
    if (threadIdx.x == 0) a.ghist[q] = 0xFFFFFFFFu;
    uint32_t* g = a.ghist + a.nq + (size_t)q * a.hist_stride;
    for (uint32_t i = threadIdx.x; i < a.hist_stride; i += TAUH_THREADS) g[i] = 0;
  }
  __shared__ uint32_t sq[32];
  if (threadIdx.x < 32) {
    const uint32_t* qw = reinterpret_cast<const uint32_t*>(a.qsigs + (size_t)q * a.w64);
    sq[threadIdx.x] = threadIdx.x < 2 * a.w64 ? (qw[threadIdx.x] & sig_word_mask(a.hash_num, threadIdx.x)) : 0u;   // padding words of a row are zero
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < S; i += TAUH_THREADS) {
    uint32_t d = 0;
    if (a.w32 == 2) {
      const uint2 v = *reinterpret_cast<const uint2*>(a.slab + (size_t)i * 8);
      d = __popc(v.x ^ sq[0]) + __popc(v.y ^ sq[1]);
    } else {
      // the row's 16-byte chunks are 512 bytes apart: issue all loads, then count
      const uint32_t C = a.w32 >> 2;
      const unsigned char* base = a.slab + ((size_t)(i >> 5) * C * 32 + (i & 31)) * 16;
      uint4 v[8];
#pragma unroll
      for (int c = 0; c < 8; ++c)
        if ((uint32_t)c < C) v[c] = *reinterpret_cast<const uint4*>(base + (size_t)c * 512);
#pragma unroll
      for (int c = 0; c < 8; ++c)
        if ((uint32_t)c < C)
          d += __popc(v[c].x ^ sq[4 * c]) + __popc(v[c].y ^ sq[4 * c + 1]) + __popc(v[c].z ^ sq[4 * c + 2]) +
               __popc(v[c].w ^ sq[4 * c + 3]);
    }
    atomicAdd(&hist[d], 1u);
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    uint32_t run = 0;
    for (uint32_t base = 0; base < bins; base += 32) {
      uint32_t v = hist[base + lane];   // bins beyond hash_num are zero
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(FULL_MASK, v, o);
        if (lane >= o) v += n;
      }
      const unsigned hit = __ballot_sync(FULL_MASK, run + v >= a.k);
      if (hit) { if (lane == 0) s_found = base + (__ffs(hit) - 1); break; }
      run += __shfl_sync(FULL_MASK, v, 31);
    }
    __syncwarp();
    if (lane == 0) {
      const uint32_t t = (a.k <= S) ? s_found : 0xFFFFFFFFu;
      tau_out[q] = (t == 0xFFFFFFFFu) ? KEY_SENTINEL : (((unsigned long long)t + 1ull) << 32);
    }
  }
}

// euclid_lsh: the sample's k-th smallest score key by an MSB-first bit search over keys held in
// registers (one barrier per varying bit) instead of a bitonic sort of the whole sample.
static const int TAUE_THREADS = 1024;   // up to 1024 * TAUE_KPT sample rows
template <int TAUE_KPT>
__global__ void __launch_bounds__(TAUE_THREADS) tau_init_euclid_kernel(ScanArgs a, uint32_t S, unsigned long long* tau_out) {
  __shared__ uint32_t sq[32];
  __shared__ uint32_t s_cnt[3];
  __shared__ uint32_t s_or, s_and;
  const uint32_t q = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31;
  if (tid < 32) {
    const uint32_t* qw = reinterpret_cast<const uint32_t*>(a.qsigs + (size_t)q * a.w64);
    sq[tid] = (uint32_t)tid < 2 * a.w64 ? (qw[tid] & sig_word_mask(a.hash_num, (uint32_t)tid)) : 0u;
  }
  if (tid == 0) { s_cnt[0] = 0; s_cnt[1] = 0; s_cnt[2] = 0; s_or = 0; s_and = 0xFFFFFFFFu; }
  if (a.ghist) {   // this query's part of the shared histogram starts from zero, its bound word from "none"
    if (tid == 0) a.ghist[q] = 0xFFFFFFFFu;
    uint32_t* g = a.ghist + a.nq + (size_t)q * a.hist_stride;
    for (uint32_t i = tid; i < a.hist_stride; i += TAUE_THREADS) g[i] = 0;
  }
  __syncthreads();
  const float two_nq_f = 2.0f * a.qnorms[q];
  uint32_t r[TAUE_KPT];
  uint32_t vor = 0, vand = 0xFFFFFFFFu;
#pragma unroll
  for (int j = 0; j < TAUE_KPT; ++j) {
    const uint32_t i = j * TAUE_THREADS + tid;
    uint32_t key = 0xFFFFFFFFu;
    if (i < S) {
      uint32_t d = 0;
      if (a.w32 == 2) {
        const uint2 v = *reinterpret_cast<const uint2*>(a.slab + (size_t)i * 8);
        d = __popc(v.x ^ sq[0]) + __popc(v.y ^ sq[1]);
      } else {
        const uint32_t C = a.w32 >> 2;
        const unsigned char* base = a.slab + ((size_t)(i >> 5) * C * 32 + (i & 31)) * 16;
        uint4 v[8];
#pragma unroll
        for (int c = 0; c < 8; ++c)
          if ((uint32_t)c < C) v[c] = *reinterpret_cast<const uint4*>(base + (size_t)c * 512);
#pragma unroll
        for (int c = 0; c < 8; ++c)
          if ((uint32_t)c < C)
            d += __popc(v[c].x ^ sq[4 * c]) + __popc(v[c].y ^ sq[4 * c + 1]) + __popc(v[c].z ^ sq[4 * c + 2]) +
                 __popc(v[c].w ^ sq[4 * c + 3]);
      }
      key = float_key(euclid_exact_score(a.norms[i], two_nq_f, a.coslut[d], a.score_mode));
      vor |= key; vand &= key;
    }
    r[j] = key;
  }
  vor = __reduce_or_sync(FULL_MASK, vor);
  vand = __reduce_and_sync(FULL_MASK, vand);
  if (lane == 0) { atomicOr(&s_or, vor); atomicAnd(&s_and, vand); }
  __syncthreads();
  uint32_t t = 0xFFFFFFFFu;
  if (a.k <= S) {
    const uint32_t all_or = s_or, all_and = s_and, varying = all_or ^ all_and;
    uint32_t prefix = 0, round = 0;
    for (int bit = 31; bit >= 0; --bit) {
      if (!((varying >> bit) & 1u)) { prefix |= all_and & (1u << bit); continue; }
      const uint32_t trial = prefix | (1u << bit);
      uint32_t c = 0;
#pragma unroll
      for (int j = 0; j < TAUE_KPT; ++j) c += (r[j] < trial) ? 1u : 0u;   // the padding keys (all ones) never count
      c = __reduce_add_sync(FULL_MASK, c);
      const uint32_t slot = round % 3u;
      if (lane == 0 && c) atomicAdd(&s_cnt[slot], c);
      if (tid == 0) s_cnt[(round + 1u) % 3u] = 0;
      __syncthreads();
      if (s_cnt[slot] < a.k) prefix = trial;
      ++round;
    }
    t = prefix;   // the k-th smallest key of the sample
  }
  if (tid == 0) tau_out[q] = (t == 0xFFFFFFFFu) ? KEY_SENTINEL : (((unsigned long long)t + 1ull) << 32);
}

template <bool EUCLID>
static cudaError_t launch_scan_w(uint32_t w32, const ScanArgs& a, const ScanConfig& cfg, dim3 grid,
                                 cudaStream_t s) {
  switch (w32) {
    case 2: return launch_scan_t<2, EUCLID, 0>(a, cfg, grid, s);
    case 4: return launch_scan_pv<4, EUCLID>(a, cfg, grid, s);
    case 8: return launch_scan_pv<8, EUCLID>(a, cfg, grid, s);
    case 16: return launch_scan_pv<16, EUCLID>(a, cfg, grid, s);
    case 32: return launch_scan_pv<32, EUCLID>(a, cfg, grid, s);
    default: return cudaErrorInvalidValue;
  }
}

cudaError_t launch_scan(const TableView& t, bool euclid, const ScanConfig& cfg, const uint64_t* d_qsigs,
                        const float* d_qnorms, const double* d_coslut, uint32_t nq, uint32_t k,
                        const uint64_t* d_floor_keys, uint64_t* d_tau_scratch, uint32_t* d_hist_scratch,
                        uint64_t* d_cand, cudaStream_t s, int* kernels_launched) {
  ScanArgs a;
  a.slab = static_cast<const unsigned char*>(t.slab);
  a.norms = t.norms;
  a.local_rows = t.local_rows;
  a.ntiles = (t.local_rows + 31) / 32;
  a.qsigs = d_qsigs; a.qnorms = d_qnorms; a.coslut = d_coslut; a.floor_keys = d_floor_keys;
  a.tau_init = nullptr;
  a.ghist = nullptr; a.hist_stride = hist_stride_for(t.hash_num, euclid);
  a.cand = d_cand;
  a.nq = nq; a.k = k; a.cap = cfg.cap; a.qp = cfg.qp; a.w64 = t.w64; a.hash_num = t.hash_num;
  a.tiles_per_chunk = cfg.tiles_per_chunk; a.nchunks = cfg.nchunks;
  a.w32 = t.w32;
  a.interleave = cfg.interleave;
  a.score_mode = cfg.score_mode;
  static const uint32_t dbg = getenv("JUBANN_SCAN_DEBUG") ? (uint32_t)atoi(getenv("JUBANN_SCAN_DEBUG")) : 0u;
  a.debug = dbg;
  a.trace = nullptr;
  static const char* trace_path = getenv("JUBANN_SCAN_TRACE");
  static unsigned long long* d_trace = nullptr;
  const uint32_t trace_groups = (nq + cfg.qp - 1) / cfg.qp;
  if (trace_path && cfg.variant == 2 && trace_groups <= 64) {
    if (!d_trace) cudaMalloc(&d_trace, 4 * 64 * 8 * 8);
    cudaMemsetAsync(d_trace, 0, 4 * 64 * 8 * 8, s);
    a.trace = d_trace;
  }
  const bool use_hist = d_hist_scratch && t.local_rows >= (1u << 16) && !(dbg & 2u);
  if (use_hist) a.ghist = d_hist_scratch;
  bool hist_cleared = false;
  // sample pre-pass (not with a pagination floor: the bound would not hold for keys above it),
  // worthwhile only when the table is much larger than the sample
  if (d_tau_scratch && !d_floor_keys) {
    uint64_t S = std::max<uint64_t>(2048, 4ull * k);
    if (S > 8192) S = 8192;
    // ranking a sample is cheap (counting / bit search): take the largest one the table allows
    if (t.local_rows >= 16 * 8192) S = 8192;
    // a launch of many passes pays for a loose bound at the start of every one of them (the first stage
    // of a pass ran at a quarter of the speed of the others on a 12.5 M-row shard): a larger sample then
    const uint32_t groups = (nq + cfg.qp - 1) / cfg.qp;
    if (groups >= 8 && t.local_rows >= (1u << 22)) S = euclid ? 16384 : 32768;
    // rows of up to 128 bits (minhash): candidates are dearer relative to a row's popcounts, and the
    // sample is cheap: 65536 rows took a 2.5 M-row shard at k=50 from 0.109 to 0.094 ms per pass
    if (groups >= 8 && !euclid && t.w32 <= 4 && t.local_rows >= (1u << 20)) S = 65536;
    // small tables: a quarter of the rows at most (without any bound a CTA's first rows all go
    // through the list, which is what a query on a 10^4-row table spent its time in)
    static const long forced_sample = getenv("JUBANN_SAMPLE_ROWS") ? atol(getenv("JUBANN_SAMPLE_ROWS")) : 0;   // tuning runs
    if (forced_sample > 0 && !euclid) S = (uint64_t)forced_sample;
    while (S > 256 && t.local_rows < 4 * S) S >>= 1;
    if (t.local_rows >= 4 * S && k <= S) {
      if (euclid) {
        if (S > 8192) tau_init_euclid_kernel<16><<<nq, TAUE_THREADS, 0, s>>>(a, (uint32_t)S, reinterpret_cast<unsigned long long*>(d_tau_scratch));
        else tau_init_euclid_kernel<8><<<nq, TAUE_THREADS, 0, s>>>(a, (uint32_t)S, reinterpret_cast<unsigned long long*>(d_tau_scratch));
        hist_cleared = use_hist;
      } else {
        tau_init_hamming_kernel<<<nq, TAUH_THREADS, 0, s>>>(a, (uint32_t)S, reinterpret_cast<unsigned long long*>(d_tau_scratch));
        hist_cleared = use_hist;
      }
      cudaError_t e = cudaGetLastError();
      if (e != cudaSuccess) return e;
      if (kernels_launched) ++*kernels_launched;
      a.tau_init = d_tau_scratch;
    }
  }
  if (use_hist && !hist_cleared) {
    cudaError_t e = cudaMemsetAsync(d_hist_scratch, 0xFF, (size_t)nq * 4, s);   // published bounds: none yet
    if (e == cudaSuccess) e = cudaMemsetAsync(d_hist_scratch + nq, 0, (size_t)nq * a.hist_stride * 4, s);
    if (e != cudaSuccess) return e;
  }
  dim3 grid(cfg.nchunks, (nq + cfg.qp - 1) / cfg.qp, 1);
  if (kernels_launched) ++*kernels_launched;
  const cudaError_t launched = euclid ? launch_scan_w<true>(t.w32, a, cfg, grid, s) : launch_scan_w<false>(t.w32, a, cfg, grid, s);
  if (a.trace && launched == cudaSuccess) {   // tuning aid: the last launch's timestamps, one line per (CTA, group)
    std::vector<unsigned long long> hbuf(4 * 64 * 8);
    cudaStreamSynchronize(s);
    cudaMemcpy(hbuf.data(), d_trace, hbuf.size() * 8, cudaMemcpyDeviceToHost);
    if (FILE* f = fopen(trace_path, "w")) {
      for (uint32_t c = 0; c < 4; ++c)
        for (uint32_t g = 0; g < trace_groups; ++g) {
          fprintf(f, "%u %u", c, g);
          for (int j = 0; j < 8; ++j) fprintf(f, " %llu", hbuf[((size_t)c * trace_groups + g) * 8 + j]);
          fprintf(f, "\n");
        }
      fclose(f);
    }
  }
  return launched;
}

cudaError_t launch_merge_local(const uint64_t* d_cand, uint32_t nq, uint32_t nlists, uint32_t k,
                               const MergeOut& out, cudaStream_t s, uint32_t* d_dense, int* kernels_launched) {
  const uint64_t M = (uint64_t)nlists * k;
  const unsigned long long* c = reinterpret_cast<const unsigned long long*>(d_cand);
  // the sparse reduction first (when the caller has nq words of scratch for the flags and there is
  // something to skip: a single short list is as quick to select from directly)
  const uint32_t* only = nullptr;
  if (d_dense && M > 2048) {
    merge_sparse_kernel<<<nq, MSP_THREADS, 0, s>>>(c, nlists, k, out, d_dense);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    if (kernels_launched) ++*kernels_launched;
    only = d_dense;
  }
  if (kernels_launched) ++*kernels_launched;
  if (M <= MSEL_MAX_KEYS) {
    uint32_t kp = 32;
    while (kp < k) kp <<= 1;
    const size_t sm = (size_t)kp * 8 * (kp <= 512 ? 2 : 1);
    if (M <= 128u * 4u) merge_small_kernel<4, 128><<<nq, 128, sm, s>>>(c, nlists, k, kp, out, only);
    else if (M <= 128u * 16u) merge_small_kernel<16, 128><<<nq, 128, sm, s>>>(c, nlists, k, kp, out, only);
    else if (M <= 128u * 32u) merge_small_kernel<32, 128><<<nq, 128, sm, s>>>(c, nlists, k, kp, out, only);
    else if (M <= 1024u * 8u) merge_small_kernel<8, 1024><<<nq, 1024, sm, s>>>(c, nlists, k, kp, out, only);
    else if (M <= 1024u * 16u) merge_small_kernel<16, 1024><<<nq, 1024, sm, s>>>(c, nlists, k, kp, out, only);
    else merge_small_kernel<24, 1024><<<nq, 1024, sm, s>>>(c, nlists, k, kp, out, only);
    return cudaGetLastError();
  }
  uint32_t half = 2 * MERGE_THREADS;
  while (half < k) half <<= 1;
  const size_t smem = (size_t)2 * half * 8;
  cudaError_t e = cudaFuncSetAttribute(merge_local_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  merge_local_kernel<<<nq, MERGE_THREADS, smem, s>>>(c, nlists, k, half, out, only);
  return cudaGetLastError();
}

cudaError_t launch_merge_parts(const uint32_t* d_cand_key, const uint64_t* d_cand_index, uint32_t nq,
                               uint32_t k, uint32_t parts, const MergeOut& out, cudaStream_t s,
                               size_t key_stride, size_t idx_stride) {
  if (key_stride == 0) key_stride = (size_t)nq * k;
  if (idx_stride == 0) idx_stride = (size_t)nq * k;
  merge_parts_kernel<<<nq, MERGE_THREADS, 0, s>>>(d_cand_key, d_cand_index, nq, k, parts, key_stride, idx_stride, out);
  return cudaGetLastError();
}

}  // namespace jbn
