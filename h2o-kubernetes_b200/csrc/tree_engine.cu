// tree_engine.cu — the resident, node-grouped tree engine (see engine.h for the design).
//
// Kernels:
//   eng_hist_kernel      one level's histogram pass: CTA = (column, chunk of one segment); each segment holds the
//                        rows of at most two nodes, whose descriptors and observed ranges live in registers; 32 private
//                        histogram copies per node (bank == lane: conflict-free, no same-address serialisation); when
//                        the next layout has been prepared every x is also stored at its new position (fused regroup).
//   eng_route_kernel     a6: row -> child node (x >= splat, NA by direction), leaves recorded, per-tile group counts.
//   eng_scan_kernel(s)   per-group exclusive scan over tiles, segment table of the next layout (32-aligned segments).
//   eng_scatter_kernel   destination position of every row; row records / weights / row ids move to the next layout.
#include "common.cuh"
#include "engine.h"
#include "kernels.h"

namespace h2ok {

namespace {

constexpr int kSegLo = 0;
constexpr int kSegHi = kEngMaxGroups;
constexpr int kSegUsed = 2 * kEngMaxGroups;  // extent of the layout (end of the dead zone)
constexpr int kSegDeadLo = kSegUsed + 1;     // dead zone: rows that already sit in a leaf
constexpr int kSegInts = kSegUsed + 4;

__global__ void eng_identity_kernel(int32_t* rowid, int32_t* seg, int64_t rows, int64_t cap) {
  int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p < cap) rowid[p] = p < rows ? (int32_t)p : -1;
  if (p == 0) {
    seg[kSegLo] = 0;
    seg[kSegHi] = (int32_t)rows;
    seg[kSegUsed] = (int32_t)rows;
    seg[kSegDeadLo] = (int32_t)rows;
  }
}

// New tree: every row goes back to the root.  The layout of the previous tree is kept (any order is as good as any
// other); its whole extent becomes one segment and padding positions stay dead.
// Engine row record (8 B, what the histogram pass reads): lo = wyq << 5 | code, hi = wyyq << 5, where code is the
// row's side within its segment for the coming pass: 0 = first child, 1 = second child, 2 = takes no part.
// The row's node id lives in a row-level array (pnode) that only the routing kernels read.
constexpr unsigned kCodeDead = 2u;

__global__ void eng_begin_kernel(const int32_t* __restrict__ rowid, const int32_t* __restrict__ wq_rows,
                                 const int32_t* __restrict__ wyq_rows, const int32_t* __restrict__ wyyq_rows,
                                 int32_t* __restrict__ seg, int32_t* __restrict__ seg_child, uint2* __restrict__ rec,
                                 int32_t* __restrict__ pnode, int32_t* __restrict__ wq, int64_t cap) {
  const int used = seg[kSegUsed];
  int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p == 0) {
    seg[kSegLo] = 0;
    seg[kSegHi] = used;
    seg[kSegDeadLo] = used;
    seg_child[0] = 0;
    seg_child[1] = -1;
  }
  if (p >= cap) return;
  uint2 r = make_uint2(kCodeDead, 0u);
  int w = 0, node = -1;
  if (p < used) {
    const int rid = rowid[p];
    if (rid >= 0) {
      w = wq_rows ? wq_rows[rid] : 1;
      r = make_uint2((unsigned)wyq_rows[rid] << 5, (unsigned)wyyq_rows[rid] << 5);  // code 0: the root
      node = 0;  // zero-weight rows still travel down the tree (a6); the histogram pass skips them by weight
    }
  }
  rec[p] = r;
  pnode[p] = node;
  if (wq) wq[p] = w;
}

// Work items: every segment (and the dead zone, which only has to be copied) cut into chunks of chunk_rows.
__global__ void __launch_bounds__(1024) eng_items_kernel(const int32_t* __restrict__ seg, int n_seg, int chunk_rows,
                                                         int max_items, int4* __restrict__ items,
                                                         int32_t* __restrict__ n_items) {
  __shared__ int s_scan[1024];
  __shared__ int s_base;
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  const int total = n_seg + 1;  // + dead zone
  for (int first = 0; first < total; first += 1024) {
    const int s = first + threadIdx.x;
    int lo = 0, hi = 0;
    if (s < n_seg) {
      lo = seg[kSegLo + s];
      hi = seg[kSegHi + s];
    } else if (s == n_seg) {
      lo = seg[kSegDeadLo];
      hi = seg[kSegUsed];
    }
    const int nch = hi > lo ? (hi - lo + chunk_rows - 1) / chunk_rows : 0;
    s_scan[threadIdx.x] = nch;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
      int t = threadIdx.x >= o ? s_scan[threadIdx.x - o] : 0;
      __syncthreads();
      s_scan[threadIdx.x] += t;
      __syncthreads();
    }
    const int off = s_base + s_scan[threadIdx.x] - nch;
    for (int j = 0; j < nch; j++) {
      if (off + j < max_items) {
        const int a = lo + j * chunk_rows;
        items[off + j] = make_int4(s, a, min(hi, a + chunk_rows), 0);
      }
    }
    __syncthreads();
    if (threadIdx.x == 1023) s_base += s_scan[1023];
    __syncthreads();
  }
  if (threadIdx.x == 0) *n_items = min(s_base, max_items);
}

// ------------------------------------------------------------------------------------------------ histogram pass
struct EngHistArgs {
  const float* X;   // current layout, column stride cap
  float* Xn;        // next layout (WRITE)
  int64_t cap;
  int cols;
  const uint2* rec;
  const int32_t* wq;
  const int32_t* dest;  // WRITE: position of each row in the next layout (-1: padding, not carried over)
  const int4* items;
  const int32_t* n_items;
  const int32_t* seg_child;
  int n_seg;
  const h2ok_col_desc* descs;
  int nb1;
  int64_t* acc;
  uint32_t* min2_enc;
  uint32_t* maxin_enc;
};

// 32-bit shared-window addressing: keeps the per-row address math to one IMAD (no generic->shared conversion)
__device__ __forceinline__ void sh_inc(unsigned addr) { asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(addr) : "memory"); }
__device__ __forceinline__ unsigned sh_add(unsigned addr, unsigned v) {
  unsigned o;
  asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(o) : "r"(addr), "r"(v) : "memory");
  return o;
}

template <bool WEIGHTED, bool TWO>
struct EngRow {
  unsigned h0, h1;  // shared address of this lane's copy of bin 0, first / second child
  double mn0, st0, mn1, st1;
  int nb0, nb1_;
  float fmn0, fmx0, fmn1, fmx1;
};

// cold: a 32-bit word wrapped.  Everything it needs is recomputed from the kernel arguments.
template <bool TWO>
__device__ __noinline__ void eng_carry(const EngHistArgs& a, int item, int col, bool s1, int b, bool ev, bool e1, int v1, bool e2) {
  const int seg = a.items[item].x;
  const int node = a.seg_child[2 * seg + (s1 ? 1 : 0)];
  if (node < 0 || a.descs[(int64_t)node * a.cols + col].nbin <= 0) return;  // untracked: its words are never flushed
  const int64_t cell = ((int64_t)node * a.cols + col) * a.nb1 + b;
  if (ev) carry_to_global(a.acc, cell, 0, 1);
  if (e1) carry_to_global(a.acc, cell, 1, v1 < 0 ? -1 : 1);
  if (e2) carry_to_global(a.acc, cell, 2, 1);
}

template <bool WEIGHTED, bool TWO>
__device__ __forceinline__ void eng_row(EngRow<WEIGHTED, TWO>& c, const EngHistArgs& a, int item, int col, float x, unsigned lo,
                                        unsigned hi, int wq) {
  if (lo & kCodeDead) return;       // takes no part at this level
  if (WEIGHTED && wq == 0) return;  // updateHisto skips zero-weight rows before anything else
  const bool s1 = TWO && (lo & 1u);
  const int b = bin_fast(x, s1 ? c.mn1 : c.mn0, s1 ? c.st1 : c.st0, s1 ? c.nb1_ : c.nb0);
  if (TWO && s1) {
    c.fmn1 = fminf(c.fmn1, x);
    c.fmx1 = fmaxf(c.fmx1, x);
  } else {
    c.fmn0 = fminf(c.fmn0, x);  // NaN-ignoring and exact
    c.fmx0 = fmaxf(c.fmx0, x);
  }
  const unsigned p = (s1 ? c.h1 : c.h0) + (unsigned)b * (32u * 12u);
  const int v1 = (int)lo >> 5;
  const unsigned v2 = hi >> 5;
  bool ev = false;
  if (WEIGHTED) {
    unsigned o0 = sh_add(p, (unsigned)wq);
    ev = o0 + (unsigned)wq < o0;
  } else {
    sh_inc(p);
  }
  const unsigned o1 = sh_add(p + 4, (unsigned)v1);
  const unsigned n1 = o1 + (unsigned)v1;
  const bool e1 = (n1 < o1) != (v1 < 0);
  const unsigned o2 = sh_add(p + 8, v2);
  const bool e2 = o2 + v2 < o2;
  if (ev | e1 | e2) eng_carry<TWO>(a, item, col, s1, b, ev, e1, v1, e2);
}

// The two-children pass keeps two register tiles, both descriptors, four running extremes and two histogram bases
// live: at 1024 threads (64 registers) ptxas rematerialises addresses on every row, so it runs 768 threads (85
// registers).  Measured alternatives (100M x 64, ms per level): 1024 x 2 tiles with spills 30; 768 x 2 tiles 15.3-16.3;
// 512 x 3 tiles (two ahead) 17.5 — fewer warps cost more issue efficiency than the deeper ring wins in latency.
template <bool TWO>
struct EngThreads {
  static constexpr int value = TWO ? 768 : 1024;
  static constexpr int ring = 2;
};

template <bool WEIGHTED, bool TWO, bool WRITE>
__global__ void __launch_bounds__(EngThreads<TWO>::value, 1) eng_hist_kernel(const __grid_constant__ EngHistArgs a) {
  constexpr int kThreads = EngThreads<TWO>::value;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int item = blockIdx.x / a.cols;
  if (item >= *a.n_items) return;
  const int col = blockIdx.x - item * a.cols;
  const int4 it = a.items[item];
  const int tid = threadIdx.x;
  const int nb1 = a.nb1;
  int c0 = -1, c1 = -1;
  if (it.x < a.n_seg) {
    c0 = a.seg_child[2 * it.x];
    c1 = TWO ? a.seg_child[2 * it.x + 1] : -1;
  }
  h2ok_col_desc d0, d1;
  d0.nbin = d1.nbin = 0;
  d0.min = d0.step = d1.min = d1.step = 0.0;
  if (c0 >= 0) d0 = a.descs[(int64_t)c0 * a.cols + col];
  if (c1 >= 0) d1 = a.descs[(int64_t)c1 * a.cols + col];
  const bool t0 = c0 >= 0 && d0.nbin > 0, t1 = c1 >= 0 && d1.nbin > 0;
  const bool hist = t0 | t1;
  if (!WRITE && !hist) return;  // nothing to accumulate and nothing to move

  unsigned* h = reinterpret_cast<unsigned*>(smem_raw);
  const int side_words = nb1 * 32 * 3;
  const int words = (TWO ? 2 : 1) * side_words;
  if (hist)
    for (int i = tid; i < words; i += kThreads) h[i] = (i % 3 == 1) ? kWyBias : 0u;
  __syncthreads();

  EngRow<WEIGHTED, TWO> c;
  c.h0 = (unsigned)__cvta_generic_to_shared(h) + (tid & 31) * 12;
  c.h1 = c.h0 + (TWO ? side_words * 4 : 0);
  // a child that is absent or does not track this column gets a harmless 1-bin descriptor: its rows (if any) land in
  // words that are never flushed, and the hot loop needs no test for it
  c.mn0 = t0 ? d0.min : 0.0; c.st0 = t0 ? d0.step : 0.0; c.nb0 = t0 ? d0.nbin : 1;
  c.mn1 = t1 ? d1.min : 0.0; c.st1 = t1 ? d1.step : 0.0; c.nb1_ = t1 ? d1.nbin : 1;
  c.fmn0 = c.fmn1 = __int_as_float(0x7f800000);
  c.fmx0 = c.fmx1 = __int_as_float(0xff800000);

  const int r0 = it.y, n = it.z - it.y;  // segment starts are multiples of 128, chunks multiples of 4096
  const float* __restrict__ x = a.X + (int64_t)col * a.cap + r0;
  float* __restrict__ xn = WRITE ? a.Xn + (int64_t)col * a.cap : nullptr;
  const uint2* __restrict__ rec = a.rec + r0;
  const int32_t* __restrict__ wq = WEIGHTED ? a.wq + r0 : nullptr;
  const int32_t* __restrict__ dst = WRITE ? a.dest + r0 : nullptr;

  // 128-bit loads: lane L of a warp holds positions 4L..4L+3 of the warp's 128-position block.  The regroup ranks the
  // rows of a block in the same (component, lane) order (eng_scatter_kernel), so store instruction j — the j-th
  // component of every lane — writes one dense run per child instead of 32 scattered words.  Blocks are padded to
  // 128 positions by the segment alignment, and padding rows carry a trash destination, so nothing here is predicated.
  constexpr int kStep = kThreads * 4;
  const int npad = (n + 127) & ~127;
  struct Tile {
    float4 x;
    uint4 ra, rb;
    int4 d, w;
  };
  auto load_tile = [&](Tile& t, int i) {
    t.x = __ldcs(reinterpret_cast<const float4*>(x + i));
    if (hist) {
      t.ra = __ldg(reinterpret_cast<const uint4*>(rec + i));
      t.rb = __ldg(reinterpret_cast<const uint4*>(rec + i + 2));
      if (WEIGHTED) t.w = __ldg(reinterpret_cast<const int4*>(wq + i));
    }
    if (WRITE) t.d = __ldg(reinterpret_cast<const int4*>(dst + i));
  };
  auto do_tile = [&](const Tile& t) {
    if (WRITE) {
      __stcs(xn + t.d.x, t.x.x);  // streaming stores: the next layout must not push the row arrays out of L2
      __stcs(xn + t.d.y, t.x.y);
      __stcs(xn + t.d.z, t.x.z);
      __stcs(xn + t.d.w, t.x.w);
    }
    if (hist) {
      eng_row(c, a, item, col, t.x.x, t.ra.x, t.ra.y, WEIGHTED ? t.w.x : 1);
      eng_row(c, a, item, col, t.x.y, t.ra.z, t.ra.w, WEIGHTED ? t.w.y : 1);
      eng_row(c, a, item, col, t.x.z, t.rb.x, t.rb.y, WEIGHTED ? t.w.z : 1);
      eng_row(c, a, item, col, t.x.w, t.rb.z, t.rb.w, WEIGHTED ? t.w.w : 1);
    }
  };
  {
    constexpr int NB = EngThreads<TWO>::ring;
    Tile t[NB];
    int i = tid * 4;
#pragma unroll
    for (int k = 0; k < NB - 1; k++)
      if (i + k * kStep < npad) load_tile(t[k], i + k * kStep);
    while (i < npad) {
#pragma unroll
      for (int k = 0; k < NB; k++) {
        if (i + (NB - 1) * kStep < npad) load_tile(t[(k + NB - 1) % NB], i + (NB - 1) * kStep);
        do_tile(t[k]);
        i += kStep;
        if (i >= npad) goto rows_done;
      }
    }
  rows_done:;
  }
  if (!hist) return;
  // observed ranges: per-thread registers -> warp -> global (encoded, integer min / max)
  {
    unsigned e0 = __reduce_min_sync(0xffffffffu, enc_f32_bits(__float_as_uint(c.fmn0)));
    unsigned e1 = __reduce_max_sync(0xffffffffu, enc_f32_bits(__float_as_uint(c.fmx0)));
    if ((tid & 31) == 0 && t0) {
      if (e0 != kEncPosInf) atomicMin(&a.min2_enc[(int64_t)c0 * a.cols + col], e0);
      if (e1 != kEncNegInf) atomicMax(&a.maxin_enc[(int64_t)c0 * a.cols + col], e1);
    }
    if (TWO) {
      e0 = __reduce_min_sync(0xffffffffu, enc_f32_bits(__float_as_uint(c.fmn1)));
      e1 = __reduce_max_sync(0xffffffffu, enc_f32_bits(__float_as_uint(c.fmx1)));
      if ((tid & 31) == 0 && t1) {
        if (e0 != kEncPosInf) atomicMin(&a.min2_enc[(int64_t)c1 * a.cols + col], e0);
        if (e1 != kEncNegInf) atomicMax(&a.maxin_enc[(int64_t)c1 * a.cols + col], e1);
      }
    }
  }
  __syncthreads();
  // flush the low words of all 32 copies (lane-rotated: conflict-free) to the 64-bit global accumulators
  const int lane = tid & 31;
  for (int i = tid; i < (TWO ? 2 : 1) * nb1; i += kThreads) {
    const int side = i >= nb1 ? 1 : 0;
    if (!(side ? t1 : t0)) continue;
    const int b = i - side * nb1;
    const int node = side ? c1 : c0;
    unsigned long long* gs = reinterpret_cast<unsigned long long*>(a.acc) + (((int64_t)node * a.cols + col) * nb1 + b) * 3;
    const unsigned* hc = h + (size_t)i * 32 * 3;
    unsigned long long s0 = 0, s1 = 0, s2 = 0;
#pragma unroll 4
    for (int k = 0; k < 32; k++) {
      const unsigned* w = hc + ((k + lane) & 31) * 3;
      s0 += w[0];
      s1 += w[1];
      s2 += w[2];
    }
    s1 -= 32ull * kWyBias;
    if (s0) atomicAdd(gs, s0);
    if (s1) atomicAdd(gs + 1, s1);
    if (s2) atomicAdd(gs + 2, s2);
  }
}

// ------------------------------------------------------------------------------------------------ a6 + regroup
// group ids: child node ids 0..G-2, dead rows G-1; newnode -2 marks padding positions (not carried over)
__global__ void __launch_bounds__(1024) eng_route_kernel(const float* __restrict__ X, int64_t cap,
                                                         const int32_t* __restrict__ seg, uint2* __restrict__ rec,
                                                         const int32_t* __restrict__ pnode,
                                                         const int32_t* __restrict__ rowid,
                                                         const h2ok_split* __restrict__ splits,
                                                         const int32_t* __restrict__ child_ids, int gid_base, int n_nodes,
                                                         int G, bool last_level, int32_t* __restrict__ newnode,
                                                         int32_t* __restrict__ leaf, int32_t* __restrict__ tile_counts) {
  extern __shared__ int s_cnt[];
  const int used = seg[kSegUsed];
  const int tile = blockIdx.x;
  if ((int64_t)tile * kEngTile >= used) return;
  for (int g = threadIdx.x; g < G; g += 1024) s_cnt[g] = 0;
  __syncthreads();
  for (int k = 0; k < kEngTile / 1024; k++) {
    const int p = tile * kEngTile + k * 1024 + threadIdx.x;
    if (p >= used) break;
    const int rid = rowid[p];
    int nn = -2;
    if (rid >= 0) {
      const int g = pnode[p];
      nn = -1;
      unsigned code = kCodeDead;
      if (g >= 0 && g < n_nodes) {
        const int col = splits[g].col;
        if (col < 0) {
          leaf[rid] = (gid_base + g) * 3;
        } else {
          const int na = splits[g].na_dir;
          const float splat = splits[g].splat;
          const float d = X[(int64_t)col * cap + p];
          int way;
          if (d != d) way = (na == H2OK_NA_LEFT) ? 0 : 1;
          else if (na == H2OK_NA_VS_REST) way = 0;
          else way = d >= splat ? 1 : 0;
          const int ch = child_ids[2 * g + way];
          if (ch < 0) leaf[rid] = (gid_base + g) * 3 + 1 + way;
          else {
            nn = ch;
            code = (unsigned)way;  // the next pass builds this segment's two children side by side
          }
        }
      }
      if (last_level) code = kCodeDead;
      rec[p].x = (rec[p].x & ~31u) | code;
      if (!last_level) atomicAdd(&s_cnt[nn < 0 ? G - 1 : nn], 1);
    }
    newnode[p] = last_level && nn >= 0 ? -1 : nn;
  }
  if (last_level) return;
  __syncthreads();
  for (int g = threadIdx.x; g < G; g += 1024) tile_counts[(int64_t)tile * G + g] = s_cnt[g];
}

// one CTA per group: exclusive prefix of the group's counts over the tiles (in place), total to group_total
__global__ void __launch_bounds__(256) eng_scan_tiles_kernel(int32_t* __restrict__ tile_counts, const int32_t* __restrict__ seg,
                                                             int G, int32_t* __restrict__ group_total) {
  __shared__ int s_scan[256];
  __shared__ int s_run;
  const int g = blockIdx.x;
  const int used = seg[kSegUsed];
  const int tiles = (used + kEngTile - 1) / kEngTile;
  if (threadIdx.x == 0) s_run = 0;
  __syncthreads();
  for (int first = 0; first < tiles; first += 256) {
    const int t = first + threadIdx.x;
    const int v = t < tiles ? tile_counts[(int64_t)t * G + g] : 0;
    s_scan[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < 256; o <<= 1) {
      int u = threadIdx.x >= o ? s_scan[threadIdx.x - o] : 0;
      __syncthreads();
      s_scan[threadIdx.x] += u;
      __syncthreads();
    }
    if (t < tiles) tile_counts[(int64_t)t * G + g] = s_run + s_scan[threadIdx.x] - v;
    __syncthreads();
    if (threadIdx.x == 255) s_run += s_scan[255];
    __syncthreads();
  }
  if (threadIdx.x == 0) group_total[g] = s_run;
}

// segment table of the next layout: children in id order, every start a multiple of 32, the dead zone last
__global__ void eng_segments_kernel(const int32_t* __restrict__ group_total, int G, int32_t* __restrict__ seg_next) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  int pos = 0;
  for (int g = 0; g < G - 1; g++) {
    seg_next[kSegLo + g] = pos;
    seg_next[kSegHi + g] = pos + group_total[g];
    pos = (pos + group_total[g] + kEngSegAlign - 1) / kEngSegAlign * kEngSegAlign;  // blocks of 128 stay whole
  }
  seg_next[kSegDeadLo] = pos;
  seg_next[kSegUsed] = pos + group_total[G - 1];
}

__global__ void __launch_bounds__(1024) eng_scatter_kernel(const int32_t* __restrict__ seg, const int32_t* __restrict__ seg_next,
                                                           const int32_t* __restrict__ tile_counts, int G,
                                                           const int32_t* __restrict__ newnode,
                                                           const uint2* __restrict__ rec, const int32_t* __restrict__ wq,
                                                           const int32_t* __restrict__ rowid, uint2* __restrict__ rec_n,
                                                           int32_t* __restrict__ pnode_n, int32_t* __restrict__ wq_n,
                                                           int32_t* __restrict__ rowid_n, int32_t* __restrict__ dest,
                                                           int trash) {
  extern __shared__ int s_cur[];
  const int used = seg[kSegUsed];
  const int tile = blockIdx.x;
  if ((int64_t)tile * kEngTile >= used) return;
  for (int g = threadIdx.x; g < G; g += 1024)
    s_cur[g] = tile_counts[(int64_t)tile * G + g] + (g < G - 1 ? seg_next[kSegLo + g] : seg_next[kSegDeadLo]);
  __syncthreads();
  // One warp per 128-position block; lane L owns positions 4L..4L+3 exactly as in the histogram pass, and the rows of
  // a block that go to the same group are ranked component by component, lane by lane.  The histogram pass's j-th
  // store instruction then writes consecutive positions.
  const unsigned lane = threadIdx.x & 31u;
  const int p0 = tile * kEngTile + (threadIdx.x >> 5) * 128 + (int)lane * 4;
#pragma unroll
  for (int j = 0; j < 4; j++) {
    const int p = p0 + j;
    const int nn = p < used ? newnode[p] : -2;
    const int grp = nn == -2 ? -1 : (nn < 0 ? G - 1 : nn);
    const unsigned peers = __match_any_sync(0xffffffffu, grp);
    int d = trash;
    if (grp >= 0) {
      const int leader = __ffs(peers) - 1;
      int base = 0;
      if ((int)lane == leader) base = atomicAdd(&s_cur[grp], __popc(peers));
      base = __shfl_sync(peers, base, leader);
      d = base + __popc(peers & ((1u << lane) - 1u));
      rec_n[d] = rec[p];
      pnode_n[d] = nn;
      rowid_n[d] = rowid[p];
      if (wq) wq_n[d] = wq[p];
    }
    if (p < used + 128) dest[p] = d;  // the padding that completes the last block gets the trash slot too
  }
}

__global__ void eng_nid_kernel(const int32_t* __restrict__ seg, const int32_t* __restrict__ pnode,
                               const int32_t* __restrict__ rowid, int32_t* __restrict__ nid_rows) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= seg[kSegUsed]) return;
  const int rid = rowid[p];
  if (rid >= 0) nid_rows[rid] = pnode[p];
}

inline unsigned nblk(int64_t n, int t) { return (unsigned)((n + t - 1) / t); }

}  // namespace

cudaError_t engine_alloc(Engine& e, int64_t rows, int cols, bool weighted) {
  engine_free(e);
  e.rows = rows;
  e.cols = cols;
  e.weighted = weighted;
  // every regroup may add up to 127 padding positions per segment; the very last position is the trash slot that
  // padding rows are "moved" to
  e.cap = (rows + kEngSegAlign - 1) / kEngSegAlign * kEngSegAlign + (int64_t)kEngSegAlign * kEngMaxGroups + 256;
  if (e.cap >= (1ll << 31)) return cudaErrorInvalidValue;
  cudaError_t err = cudaSuccess;
  auto A = [&](void** p, size_t bytes) {
    if (err == cudaSuccess) err = cudaMalloc(p, bytes);
  };
  for (int k = 0; k < 2; k++) {
    A((void**)&e.X[k], sizeof(float) * (size_t)e.cap * cols);
    A((void**)&e.rec[k], sizeof(uint2) * (size_t)e.cap);
    A((void**)&e.rowid[k], sizeof(int32_t) * (size_t)e.cap);
    A((void**)&e.pnode[k], sizeof(int32_t) * (size_t)e.cap);
    A((void**)&e.seg[k], sizeof(int32_t) * kSegInts);
    if (weighted) A((void**)&e.wq[k], sizeof(int32_t) * (size_t)e.cap);
  }
  A((void**)&e.seg_child, sizeof(int32_t) * 2 * kEngMaxGroups);
  A((void**)&e.newnode, sizeof(int32_t) * (size_t)e.cap);
  A((void**)&e.dest, sizeof(int32_t) * (size_t)e.cap);
  A((void**)&e.leaf, sizeof(int32_t) * (size_t)(rows > 0 ? rows : 1));
  e.tile_cap = (e.cap + kEngTile - 1) / kEngTile;
  A((void**)&e.tile_counts, sizeof(int32_t) * (size_t)(e.tile_cap + 1) * kEngMaxGroups);
  // chunk: about 16 CTAs per SM over the whole level, never below 64 Ki rows
  int64_t chunk = (rows * cols) / (16 * kSMs) / 4096 * 4096;
  if (chunk < 65536) chunk = 65536;
  e.chunk_rows = chunk;
  e.max_items = (int)(e.cap / chunk) + kEngMaxGroups + 2;
  A((void**)&e.items, sizeof(int4) * (size_t)e.max_items);
  A((void**)&e.n_items, sizeof(int32_t));
  if (err != cudaSuccess) engine_free(e);
  return err;
}

void engine_free(Engine& e) {
  for (int k = 0; k < 2; k++) {
    cudaFree(e.X[k]); cudaFree(e.rec[k]); cudaFree(e.rowid[k]); cudaFree(e.pnode[k]); cudaFree(e.seg[k]); cudaFree(e.wq[k]);
    e.X[k] = nullptr; e.rec[k] = nullptr; e.rowid[k] = nullptr; e.pnode[k] = nullptr; e.seg[k] = nullptr; e.wq[k] = nullptr;
  }
  cudaFree(e.seg_child); cudaFree(e.newnode); cudaFree(e.dest); cudaFree(e.leaf); cudaFree(e.tile_counts);
  cudaFree(e.items); cudaFree(e.n_items);
  e.seg_child = e.newnode = e.dest = e.leaf = e.tile_counts = e.n_items = nullptr;
  e.items = nullptr;
  e.rows = e.cap = 0;
}

cudaError_t engine_reset_layout(Engine& e, cudaStream_t s) {
  e.cur = 0;
  e.n_seg = 1;
  e.dest_valid = false;
  eng_identity_kernel<<<nblk(e.cap, 256), 256, 0, s>>>(e.rowid[0], e.seg[0], e.rows, e.cap);
  return cudaGetLastError();
}

cudaError_t engine_begin_tree(Engine& e, const int32_t* wq_rows, const int32_t* wyq_rows, const int32_t* wyyq_rows,
                              cudaStream_t s) {
  cudaError_t err = cudaMemsetAsync(e.leaf, 0xFF, sizeof(int32_t) * (size_t)e.rows, s);
  if (err != cudaSuccess) return err;
  eng_begin_kernel<<<nblk(e.cap, 256), 256, 0, s>>>(e.rowid[e.cur], e.weighted ? wq_rows : nullptr, wyq_rows, wyyq_rows,
                                                    e.seg[e.cur], e.seg_child, e.rec[e.cur], e.pnode[e.cur],
                                                    e.weighted ? e.wq[e.cur] : nullptr, e.cap);
  e.n_seg = 1;
  e.dest_valid = false;
  return cudaGetLastError();
}

template <bool W, bool TWO, bool WR>
static cudaError_t launch_eng_hist(const EngHistArgs& a, unsigned grid, size_t smem, cudaStream_t s) {
  static bool attr_done[64] = {};  // per device, per instantiation
  int dev = 0;
  cudaGetDevice(&dev);
  bool& attr = attr_done[dev & 63];
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(eng_hist_kernel<W, TWO, WR>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    attr = true;
  }
  eng_hist_kernel<W, TWO, WR><<<grid, EngThreads<TWO>::value, smem, s>>>(a);
  return cudaGetLastError();
}

cudaError_t engine_build_level(Engine& e, int n_nodes, const h2ok_col_desc* descs, int nb1, int64_t* acc,
                               uint32_t* min2_enc, uint32_t* maxin_enc, cudaStream_t s) {
  (void)n_nodes;
  const bool two = e.dest_valid;  // the root pass has one node per segment, every later pass the two children
  const bool write = e.dest_valid;
  if ((size_t)(two ? 2 : 1) * nb1 * 32 * 12 > 227 * 1024) return cudaErrorInvalidValue;
  if (e.chunk_rows % kEngTile) return cudaErrorInvalidValue;
  eng_items_kernel<<<1, 1024, 0, s>>>(e.seg[e.cur], e.n_seg, (int)e.chunk_rows, e.max_items, e.items, e.n_items);
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return err;
  EngHistArgs a{};
  a.X = e.X[e.cur];
  a.Xn = e.X[1 - e.cur];
  a.cap = e.cap;
  a.cols = e.cols;
  a.rec = e.rec[e.cur];
  a.wq = e.weighted ? e.wq[e.cur] : nullptr;
  a.dest = e.dest;
  a.items = e.items;
  a.n_items = e.n_items;
  a.seg_child = e.seg_child;
  a.n_seg = e.n_seg;
  a.descs = descs;
  a.nb1 = nb1;
  a.acc = acc;
  a.min2_enc = min2_enc;
  a.maxin_enc = maxin_enc;
  const int items_bound = (int)(e.cap / e.chunk_rows) + e.n_seg + 2;
  const unsigned grid = (unsigned)e.cols * (unsigned)(items_bound < e.max_items ? items_bound : e.max_items);
  const size_t smem = (size_t)(two ? 2 : 1) * nb1 * 32 * 12;
  if (e.weighted) {
    if (two) err = launch_eng_hist<true, true, true>(a, grid, smem, s);
    else err = launch_eng_hist<true, false, false>(a, grid, smem, s);
  } else {
    if (two) err = launch_eng_hist<false, true, true>(a, grid, smem, s);
    else err = launch_eng_hist<false, false, false>(a, grid, smem, s);
  }
  if (err != cudaSuccess) return err;
  if (write) {  // the next layout is complete: it becomes the current one
    e.cur = 1 - e.cur;
    e.dest_valid = false;
  }
  return cudaSuccess;
}

cudaError_t engine_apply_splits(Engine& e, const h2ok_split* splits, const int32_t* child_ids, int gid_base, int n_nodes,
                                bool last_level, cudaStream_t s) {
  const int G = 2 * n_nodes + 1;
  if (G > kEngMaxGroups) return cudaErrorInvalidValue;
  const unsigned tiles = (unsigned)e.tile_cap;
  int32_t* group_total = e.tile_counts + (size_t)e.tile_cap * kEngMaxGroups;
  eng_route_kernel<<<tiles, 1024, sizeof(int) * G, s>>>(e.X[e.cur], e.cap, e.seg[e.cur], e.rec[e.cur], e.pnode[e.cur],
                                                        e.rowid[e.cur], splits, child_ids, gid_base, n_nodes, G, last_level,
                                                        e.newnode, e.leaf, e.tile_counts);
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return err;
  if (last_level)  // every row now sits in a leaf
    return cudaMemsetAsync(e.pnode[e.cur], 0xFF, sizeof(int32_t) * (size_t)e.cap, s);
  const int nx = 1 - e.cur;
  eng_scan_tiles_kernel<<<G, 256, 0, s>>>(e.tile_counts, e.seg[e.cur], G, group_total);
  eng_segments_kernel<<<1, 32, 0, s>>>(group_total, G, e.seg[nx]);
  err = cudaMemsetAsync(e.rec[nx], 0xFF, sizeof(uint2) * (size_t)e.cap, s);  // node field 0x3FF everywhere: dead
  if (err == cudaSuccess) err = cudaMemsetAsync(e.rowid[nx], 0xFF, sizeof(int32_t) * (size_t)e.cap, s);
  if (err == cudaSuccess) err = cudaMemsetAsync(e.pnode[nx], 0xFF, sizeof(int32_t) * (size_t)e.cap, s);
  if (err != cudaSuccess) return err;
  eng_scatter_kernel<<<tiles, 1024, sizeof(int) * G, s>>>(e.seg[e.cur], e.seg[nx], e.tile_counts, G, e.newnode, e.rec[e.cur],
                                                          e.weighted ? e.wq[e.cur] : nullptr, e.rowid[e.cur], e.rec[nx],
                                                          e.pnode[nx], e.weighted ? e.wq[nx] : nullptr, e.rowid[nx], e.dest,
                                                          (int)(e.cap - 1));
  err = cudaGetLastError();
  if (err != cudaSuccess) return err;
  // segment s of the CURRENT layout holds node s of this level, whose children the next pass builds
  err = cudaMemcpyAsync(e.seg_child, child_ids, sizeof(int32_t) * 2 * (size_t)n_nodes, cudaMemcpyDeviceToDevice, s);
  if (err != cudaSuccess) return err;
  e.n_seg = n_nodes;
  e.dest_valid = true;
  return cudaSuccess;
}

cudaError_t engine_nid(Engine& e, int32_t* nid_rows, cudaStream_t s) {
  cudaError_t err = cudaMemsetAsync(nid_rows, 0xFF, sizeof(int32_t) * (size_t)e.rows, s);
  if (err != cudaSuccess) return err;
  eng_nid_kernel<<<nblk(e.cap, 256), 256, 0, s>>>(e.seg[e.cur], e.pnode[e.cur], e.rowid[e.cur], nid_rows);
  return cudaGetLastError();
}

}  // namespace h2ok
