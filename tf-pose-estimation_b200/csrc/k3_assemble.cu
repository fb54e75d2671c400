// k3_assemble.cu -- K3: one block per frame does what pafprocess.cpp:96-189 does after the
// candidates exist:
//   1. per-limb ordering of the candidates (std::sort with comp_candidate, :97,:239-241) --
//      one bitonic sort of the frame's 64-bit keys (limb | ~score | idx1 | idx2), i.e. score
//      descending with (idx1, idx2) ascending as the tie order (what libstdc++ gives for <= 16
//      candidates in a-major/b-minor input order when scores differ; bit-equal scores raise
//      TFPD_FLAG_SCORE_TIE because the reference's order is then introsort-internal),
//   2. greedy matching (:98-124): one warp per limb walks its sorted segment,
//   3. subset assembly (:127-183) by warp 0, the row scan spread over the lanes,
//      in `asbuilt` (found==2 never merges: `int membership;` UB at :152 as compiled by GCC -O2)
//      or `upstream` (membership = 0) mode,
//   4. the filter (:185-189) and the getter view (:194-214) -> tfpd_human records.
#include <type_traits>

#include <cstdlib>

#include "tfpd_internal.cuh"
#include "std_sort_literal.h"

namespace tfpd {

namespace {

constexpr int kWarpSortMax = 1024;           // longest limb segment a single warp sorts
constexpr int kRegRowsMax = 128;            // subset rows held in registers (4 per lane) before spilling to smem

constexpr int kK3Threads = 640;             // 20 warps: 19 greedy walkers + sort helpers
constexpr int kBitmapWords = kMaxPeaksHard / 32;   // 128

// bitonic network with a mirrored first stage per block, one warp, ascending: every comparator puts the smaller key at
// the lower index, so the (virtual) +inf padding beyond the segment never moves and is simply skipped
__device__ void warp_bitonic(unsigned long long *seg, int len, int lane) {
    int np2 = 1;
    while (np2 < len) np2 <<= 1;
    for (int k = 2; k <= np2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            const int mask = (j == (k >> 1)) ? (k - 1) : j;
            for (int i = lane; i < len; i += 32) {
                const int l = i ^ mask;
                if (l > i && l < len) {
                    const unsigned long long a = seg[i], b = seg[l];
                    if (a > b) {
                        seg[i] = b;
                        seg[l] = a;
                    }
                }
            }
            __syncwarp();
        }
    }
}

// MINB = CTAs per SM the register allocation allows: 1 -> 78 registers, 2 -> 48 registers (no spills either way).
// The kernel is a latency-bound chain per frame (issue slots ~27 % busy with one CTA per SM), so two co-resident
// frames halve the SM-time of a batch whenever their shared memory fits -- which matters when the kernel only gets
// the SMs a reduced map-kernel grid leaves free (bench.py's pipelined region) and on crowded frames.
template <int MINB>
__global__ void __launch_bounds__(kK3Threads, MINB) k_assemble(Workspace ws, Thresholds th) {
    extern __shared__ __align__(16) unsigned char sm_raw[];
    // (TFPD_GUARD_BYTES is 0 in the product build; the checked build puts a guard band behind every sub-array)
    unsigned long long *keys = reinterpret_cast<unsigned long long *>(sm_raw);            // [max_cands]
    unsigned char *g_keys = reinterpret_cast<unsigned char *>(keys + ws.max_cands);
    float *rows = reinterpret_cast<float *>(g_keys + TFPD_GUARD_BYTES);                  // [20][max_rows] SoA
    unsigned char *g_rows = reinterpret_cast<unsigned char *>(rows + 20 * (size_t)ws.max_rows);
    uint32_t *used = reinterpret_cast<uint32_t *>(g_rows + TFPD_GUARD_BYTES);            // [19][2][128]
    unsigned char *g_used = reinterpret_cast<unsigned char *>(used + kLimbs * 2 * kBitmapWords);
    // the frame's peak table, staged once: assembly and the getter view index it by cid (a dependent global
    // load per connection was the latency chain of this kernel)
    float *s_pscore = reinterpret_cast<float *>(g_used + TFPD_GUARD_BYTES);              // [max_peaks]
    int16_t *s_pkx = reinterpret_cast<int16_t *>(s_pscore + ws.max_peaks);               // [max_peaks]
    int16_t *s_pky = s_pkx + ws.max_peaks;                                               // [max_peaks]
    unsigned char *g_end = reinterpret_cast<unsigned char *>(s_pky + ws.max_peaks);
    (void)g_end;
    __shared__ int s_lstart[kLimbs + 1];
    __shared__ int s_ccount[kLimbs];
    __shared__ int s_first[kHeatC + 1];
    __shared__ int s_flags;

    __shared__ int s_lcount[kLimbs], s_lfill[kLimbs];
    __shared__ int s_maxseg;
    __shared__ float s_stage[3][kRegRowsMax];   // new subset rows on their way from connection lanes to row lanes

    const int n = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
#ifdef TFPD_CHECKED
    guard_fill(g_keys, tid, kK3Threads);
    guard_fill(g_rows, tid, kK3Threads);
    guard_fill(g_used, tid, kK3Threads);
    guard_fill(g_end, tid, kK3Threads);
#endif
    if (ws.dbg_times && tid == 0 && n < 64) ws.dbg_times[2048 + n * 8 + 0] = clock64();
    const int MR = ws.max_rows;
    const int raw = ws.counters[n].n_cand;
    const int cnt = min(raw, ws.max_cands);
    const unsigned long long *gkeys = ws.cand_key + (size_t)n * ws.max_cands;
    for (int i = tid; i < kLimbs * 2 * kBitmapWords; i += kK3Threads) used[i] = 0u;
    if (tid <= kHeatC) s_first[tid] = ws.part_first[n * 20 + tid];
    if (tid < kLimbs) s_ccount[tid] = s_lcount[tid] = s_lfill[tid] = 0;
    if (tid == 0) s_flags = (raw > ws.max_cands) ? TFPD_FLAG_CAND_OVERFLOW : 0;
    {
        const int npk = min(ws.part_first[n * 20 + kParts], ws.max_peaks);
        const size_t po = (size_t)n * ws.max_peaks;
        for (int i = tid; i < npk; i += kK3Threads) {
            s_pscore[i] = ws.pk_score[po + i];
            s_pkx[i] = ws.pk_x[po + i];
            s_pky[i] = ws.pk_y[po + i];
        }
    }
    __syncthreads();
    if (ws.dbg_times && tid == 0 && n < 64) ws.dbg_times[2048 + n * 8 + 1] = clock64();

    // ---- 1. per-limb ordering ----
    // The key's top byte is the limb, so the frame-wide order is: limb segments, each sorted.  Bucket the
    // candidates by limb (count, prefix, scatter: two passes over the global list), then every limb's segment
    // is sorted by its own warp with only warp-level synchronisation -- a frame-wide bitonic sort pays a
    // 640-thread barrier per stage (66 stages for 2048 candidates).
    for (int i = tid; i < cnt; i += kK3Threads) atomicAdd(&s_lcount[(int)(gkeys[i] >> 56)], 1);
    __syncthreads();
    if (tid == 0) {
        int acc = 0, mx = 0;
        for (int l = 0; l < kLimbs; ++l) {
            s_lstart[l] = acc;
            acc += s_lcount[l];
            mx = max(mx, s_lcount[l]);
        }
        s_lstart[kLimbs] = acc;
        s_maxseg = mx;
    }
    __syncthreads();
    for (int i = tid; i < cnt; i += kK3Threads) {
        const unsigned long long key = gkeys[i];
        const int limb = (int)(key >> 56);
        keys[s_lstart[limb] + atomicAdd(&s_lfill[limb], 1)] = key;
    }
    __syncthreads();
    if (ws.dbg_times && tid == 0 && n < 64) ws.dbg_times[2048 + n * 8 + 2] = clock64();
    if (s_maxseg <= min(kWarpSortMax, ws.k3_warp_sort_max)) {
        if (warp < kLimbs) warp_bitonic(keys + s_lstart[warp], s_lcount[warp], lane);
    } else {
        // a very long segment (one limb with > kWarpSortMax candidates): frame-wide bitonic sort by all threads;
        // the key order (limb first) keeps the segments where they are
        int np2 = 32;
        while (np2 < cnt) np2 <<= 1;
        for (int i = cnt + tid; i < np2; i += kK3Threads) keys[i] = ~0ull;
        __syncthreads();
        for (int k = 2; k <= np2; k <<= 1) {
            for (int j = k >> 1; j > 0; j >>= 1) {
                for (int i = tid; i < np2; i += kK3Threads) {
                    int ixj = i ^ j;
                    if (ixj > i) {
                        bool up = ((i & k) == 0);
                        unsigned long long a = keys[i], b = keys[ixj];
                        if ((a > b) == up) {
                            keys[i] = b;
                            keys[ixj] = a;
                        }
                    }
                }
                __syncthreads();
            }
        }
    }
    __syncthreads();
    if (ws.dbg_times && tid == 0 && n < 64) ws.dbg_times[2048 + n * 8 + 3] = clock64();
    // exact-tie detection: equal (limb, score) neighbours
    for (int i = tid + 1; i < cnt; i += kK3Threads)
        if ((keys[i] >> 24) == (keys[i - 1] >> 24)) atomicOr(&s_flags, TFPD_FLAG_SCORE_TIE);
    __syncthreads();
    if (ws.k3_std_ties && (s_flags & TFPD_FLAG_SCORE_TIE)) {
        // Rare path.  A limb segment with bit-equal scores is put back into the order the reference pushed its
        // candidates in (idx1, idx2 ascending = the key's low 24 bits, unique within a limb: the same warp network on
        // the key rotated by 24 bits) and then sorted by the literal std::sort above: the reference's own order.
        if (warp < kLimbs) {
            unsigned long long *seg = keys + s_lstart[warp];
            const int len = s_lcount[warp];
            bool tie = false;
            for (int i = lane + 1; i < len; i += 32) tie |= (seg[i] >> 24) == (seg[i - 1] >> 24);
            if (__any_sync(0xffffffffu, tie)) {
                for (int i = lane; i < len; i += 32) seg[i] = (seg[i] << 40) | (seg[i] >> 24);
                __syncwarp();
                warp_bitonic(seg, len, lane);
                for (int i = lane; i < len; i += 32) seg[i] = (seg[i] >> 40) | (seg[i] << 24);
                __syncwarp();
                if (lane == 0) std_sort_literal(seg, len);
                __syncwarp();
            }
        }
        __syncthreads();
    }

    // ---- 2. greedy matching (:98-124), one warp per limb ----
    // The walk over the sorted candidates is serial only in the ACCEPTED ones: a window of 32 candidates is
    // loaded one per lane, lanes whose end points are already taken (bitmaps, from earlier windows) drop out at
    // once, and the warp then steps from one surviving candidate to the next in order -- the accepted candidate's
    // end points are broadcast and every later lane sharing one of them drops out.  Same result as testing the
    // candidates one by one, in (number of connections) steps instead of (number of candidates).
    int16_t *c1 = ws.conn_cid1 + (size_t)n * ws.max_cands;
    int16_t *c2 = ws.conn_cid2 + (size_t)n * ws.max_cands;
    float *cs = ws.conn_score + (size_t)n * ws.max_cands;
    if (warp < kLimbs) {
        const int limb = warp;
        const int beg = s_lstart[limb], end = s_lstart[limb + 1];
        uint32_t *ua = used + (size_t)limb * 2 * kBitmapWords, *ub = ua + kBitmapWords;
        const int fa = s_first[c_limb_part[limb][0]], fb = s_first[c_limb_part[limb][1]];
        int nconn = 0;
        for (int w0 = beg; w0 < end; w0 += 32) {
            const int i = w0 + lane;
            unsigned long long key = 0ull;
            int a = -1, b = -1;
            bool alive = false;
            if (i < end) {
                key = keys[i];
                a = (int)((key >> 12) & 0xfff);
                b = (int)(key & 0xfff);
                alive = !((ua[a >> 5] >> (a & 31)) & 1u) && !((ub[b >> 5] >> (b & 31)) & 1u);   // :104-114
            }
            __syncwarp();           // every key of the window is in a register before records are packed over them
            unsigned live = __ballot_sync(0xffffffffu, alive);
            while (live) {
                const int j = __ffs(live) - 1;
                const int ja = __shfl_sync(0xffffffffu, a, j), jb = __shfl_sync(0xffffffffu, b, j);
                if (lane == j) {
                    ua[a >> 5] |= 1u << (a & 31);
                    ub[b >> 5] |= 1u << (b & 31);
                    const uint32_t sbits = 0xffffffffu - (uint32_t)((key >> 24) & 0xffffffffu);
                    // the connection record (score | cid1 | cid2; conn.cid = peak list id, :122), packed over a
                    // key this walk has already consumed (beg + nconn <= w0 + j, the window is in registers)
                    keys[beg + nconn] = ((unsigned long long)sbits << 32) |
                                        ((unsigned long long)(uint32_t)(fa + a) << 16) |
                                        (unsigned long long)(uint32_t)(fb + b);
                }
                ++nconn;
                alive = alive && lane > j && a != ja && b != jb;
                live = __ballot_sync(0xffffffffu, alive);
            }
            __syncwarp();           // bitmap updates visible to the next window
        }
        if (lane == 0) s_ccount[limb] = nconn;
        // the global connection table (tfpd_fetch_connections) in one coalesced sweep, off the serial chain
        for (int k = lane; k < nconn; k += 32) {
            const unsigned long long rec = keys[beg + k];
            c1[beg + k] = (int16_t)((rec >> 16) & 0xffffu);
            c2[beg + k] = (int16_t)(rec & 0xffffu);
            cs[beg + k] = __uint_as_float((uint32_t)(rec >> 32));
        }
    }
    __syncthreads();
    if (ws.dbg_times && tid == 0 && n < 64) ws.dbg_times[2048 + n * 8 + 4] = clock64();
    if (tid < kLimbs) {
        ws.conn_start[n * 20 + tid] = s_lstart[tid];
        ws.conn_count[n * 20 + tid] = s_ccount[tid];
    }
#ifdef TFPD_CHECKED
    if (ws.dbg_trip && tid == 0) keys[ws.max_cands] = 1ull;      // negative control: one element past the key array
    __syncthreads();
    if (lane == 0) TFPD_CHK(guard_ok(g_keys) && guard_ok(g_rows) && guard_ok(g_used) && guard_ok(g_end));
#endif
    if (warp != 0) return;

    // ---- 3. subset assembly (warp 0) ----
    // Fast path for the first kRegRowsMax rows: row r belongs to lane r & 31, slot r >> 5.  A limb only ever
    // looks at two columns of a row (its parts PA, PB) and updates PB, the score and the count, so per limb each
    // lane pulls its rows' PA / PB entries out of the smem table ONCE, keeps score and count in registers for the
    // whole assembly, and a connection is then two compares per lane and slot, a ballot and an update inside the
    // owning lane -- no shared-memory round trip on the serial chain (the all-smem loop below spends ~900 cycles
    // per connection on them) and, unlike a per-limb instantiation with statically indexed row registers, one
    // small loop body that stays in the instruction cache.  When row kRegRowsMax+1 is needed, the registers go
    // back to the smem rows and the generic loop below continues from the same (limb, connection).
    const float *pscore = s_pscore;
    const unsigned FULL = 0xffffffffu;
    int nrows = 0;
    bool row_overflow = false;
    int limb0 = kLimbs, k0 = 0;                       // where the generic loop resumes (kLimbs = nowhere)
    {
        constexpr int S = kRegRowsMax / 32;
        const int reg_cap = min(min(ws.k3_reg_rows, kRegRowsMax), MR);
        // a row that does not exist yet reads as "no part anywhere": creating one (:172-180) is then 4 writes
        for (int c = 0; c < kParts; ++c)
#pragma unroll
            for (int q = 0; q < S; ++q)
                if (lane + 32 * q < MR) rows[(size_t)c * MR + lane + 32 * q] = -1.0f;
        __syncwarp();
        float sc[S], cn[S];                                          // columns 18 / 19 of this lane's rows
        bool own[S];
#pragma unroll
        for (int q = 0; q < S; ++q) {
            sc[q] = cn[q] = 0.f;
            own[q] = lane + 32 * q < MR;
        }
        // found == 2 in upstream mode on the smem rows (:153-165); true when the two rows were merged
        auto merge_rows = [&](int r1, int r2, float cscore) -> bool {
            float *col18 = rows + 18 * (size_t)MR, *col19 = rows + 19 * (size_t)MR;
            const bool both = (lane < kParts) && (rows[(size_t)lane * MR + r1] > 0.0f) &&
                              (rows[(size_t)lane * MR + r2] > 0.0f);
            if (__ballot_sync(FULL, both) != 0u) return false;
            if (lane < kParts) {
                float *cc = rows + (size_t)lane * MR;
                cc[r1] = __fadd_rn(cc[r1], __fadd_rn(cc[r2], 1.0f));
            } else if (lane == 18) {
                col18[r1] = __fadd_rn(__fadd_rn(col18[r1], col18[r2]), cscore);
            } else if (lane == 19) {
                col19[r1] = __fadd_rn(col19[r1], col19[r2]);
            }
            __syncwarp();
            // subset.erase(begin + r2); the vacated last row reads as "not there" again
            for (int c = 0; c < 20; ++c) {
                float *cc = rows + (size_t)c * MR;
                float v[S];
#pragma unroll
                for (int q = 0; q < S; ++q) {
                    const int r = lane + 32 * q;
                    v[q] = 0.f;
                    if (r >= r2 && r < nrows) v[q] = (r + 1 < nrows) ? cc[r + 1] : -1.0f;
                }
                __syncwarp();
#pragma unroll
                for (int q = 0; q < S; ++q) {
                    const int r = lane + 32 * q;
                    if (r >= r2 && r < nrows) cc[r] = v[q];
                }
            }
            --nrows;
            __syncwarp();
            return true;
        };
        for (int limb = 0; limb < kLimbs && limb0 == kLimbs; ++limb) {
            const int pa = c_limb_part[limb][0], pb = c_limb_part[limb][1];
            float *colA = rows + (size_t)pa * MR, *colB = rows + (size_t)pb * MR;
            float av[S], bv[S];
#pragma unroll
            for (int q = 0; q < S; ++q) {
                av[q] = own[q] ? colA[lane + 32 * q] : -1.0f;
                bv[q] = own[q] ? colB[lane + 32 * q] : -1.0f;
            }
            const int beg = s_lstart[limb], nconn = s_ccount[limb];
            // one connection; NS = slots that can hold a row right now (instantiated for 1, 2 and S so that the
            // common case -- at most 32 rows -- runs the shortest instruction chain).  Returns true to bail out.
            auto step = [&](auto ns_tag, int k) -> bool {
                constexpr int NS = decltype(ns_tag)::value;
                const unsigned long long rec = keys[beg + k];
                const int cid1 = (int)((rec >> 16) & 0xffffu), cid2 = (int)(rec & 0xffffu);
                const float cscore = __uint_as_float((uint32_t)(rec >> 32));
                const float f1 = (float)cid1, f2 = (float)cid2;
                const float ps1 = pscore[cid1], ps2 = pscore[cid2];
                unsigned m[NS];
                int found = 0;
#pragma unroll
                for (int q = 0; q < NS; ++q) {                                                          // :137
                    m[q] = __ballot_sync(FULL, (lane + 32 * q < nrows) && (av[q] == f1 || bv[q] == f2));
                    found += __popc(m[q]);
                }
                if (found == 0) {
                    if (limb < 17) {                                                            // :171-181
                        if (nrows >= reg_cap) {
                            if (nrows >= MR) {
                                row_overflow = true;
                                return false;
                            }
                            limb0 = limb;               // continue in shared memory from this connection
                            k0 = k;
                            return true;
                        }
                        const float scn = __fadd_rn(__fadd_rn(ps1, ps2), cscore);
                        // (selects, not indexed stores: the row arrays must stay in registers)
                        const bool mine = lane == (nrows & 31);
#pragma unroll
                        for (int q = 0; q < S; ++q) {
                            const bool wr = mine && (nrows >= 32 * q) && (nrows < 32 * q + 32);
                            av[q] = wr ? f1 : av[q];
                            bv[q] = wr ? f2 : bv[q];
                            sc[q] = wr ? scn : sc[q];
                            cn[q] = wr ? 2.0f : cn[q];
                        }
                        ++nrows;
                    }
                } else if (found <= 2) {
                    // first (and second) matching row in row order
                    int r1 = -1, r2 = -1;
#pragma unroll
                    for (int q = 0; q < NS; ++q) {
                        unsigned mm = m[q];
                        if (mm && r1 < 0) {
                            r1 = 32 * q + __ffs(mm) - 1;
                            mm &= mm - 1u;
                        }
                        if (mm && r1 >= 0 && r2 < 0) r2 = 32 * q + __ffs(mm) - 1;
                    }
                    bool merged = false;
                    if (found == 2 && th.mode == TFPD_MODE_UPSTREAM) {
                        // :151-165, rare: done on the smem rows (one copy of the code), registers re-read after
                        float *col18 = rows + 18 * (size_t)MR, *col19 = rows + 19 * (size_t)MR;
#pragma unroll
                        for (int q = 0; q < S; ++q)
                            if (lane + 32 * q < nrows) {
                                colA[lane + 32 * q] = av[q];
                                colB[lane + 32 * q] = bv[q];
                                col18[lane + 32 * q] = sc[q];
                                col19[lane + 32 * q] = cn[q];
                            }
                        __syncwarp();
                        merged = merge_rows(r1, r2, cscore);
                        if (merged) {
#pragma unroll
                            for (int q = 0; q < S; ++q) {
                                av[q] = own[q] ? colA[lane + 32 * q] : -1.0f;
                                bv[q] = own[q] ? colB[lane + 32 * q] : -1.0f;
                                sc[q] = own[q] ? col18[lane + 32 * q] : 0.f;
                                cn[q] = own[q] ? col19[lane + 32 * q] : 0.f;
                            }
                        }
                    }
                    if (!merged) {                                                // :144-149 / :166-170
                        const bool mine = lane == (r1 & 31);
                        const float add = __fadd_rn(ps2, cscore);
#pragma unroll
                        for (int q = 0; q < NS; ++q) {
                            const bool wr = mine && (r1 >= 32 * q) && (r1 < 32 * q + 32) && (found == 2 || bv[q] != f2);
                            bv[q] = wr ? f2 : bv[q];
                            cn[q] = wr ? __fadd_rn(cn[q], 1.0f) : cn[q];
                            sc[q] = wr ? __fadd_rn(sc[q], add) : sc[q];
                        }
                    }
                } else {
                    if (lane == 0) atomicOr(&s_flags, TFPD_FLAG_FOUND_GT2);
                }
                return false;
            };
            // Whole-limb step.  Within one limb the greedy matching made every cid1 and every cid2 unique, so
            // if, in the state at the start of a pass of <= 32 connections, (1) no row holds any of the pass's
            // cid2 at PB and (2) no cid1 is held by more than one row at PA, then the connections cannot see each
            // other's effects: found is 0 or 1 for each, a matched row is matched by exactly one connection and
            // takes its update, and the unmatched ones append rows in connection order -- exactly what the
            // sequential loop (:134-183) would do, without its serial chain.  Limbs 0..16 normally qualify (their PB
            // column is still empty); anything else, and the redundant ear-shoulder limbs 17/18, go one by one.
            int kpar = 0;
            // one pass of <= 32 connections over NS row slots; false = conditions not met, nothing changed
            auto limb_pass = [&](auto ns_tag, int np) -> bool {
                constexpr int NS = decltype(ns_tag)::value;
                float cf1 = -2.0f, cf2 = -2.0f, cadd = 0.f, cnew = 0.f;
                if (lane < np) {
                    const unsigned long long rec = keys[beg + kpar + lane];
                    const int cid1 = (int)((rec >> 16) & 0xffffu), cid2 = (int)(rec & 0xffffu);
                    const float cscore = __uint_as_float((uint32_t)(rec >> 32));
                    cf1 = (float)cid1;
                    cf2 = (float)cid2;
                    cadd = __fadd_rn(pscore[cid2], cscore);                                     // :147 / :169
                    cnew = __fadd_rn(__fadd_rn(pscore[cid1], pscore[cid2]), cscore);            // :176-178
                }
                unsigned mybits[NS];
                bool anyb = false;
#pragma unroll
                for (int q = 0; q < NS; ++q) mybits[q] = 0u;
                for (int j = 0; j < np; ++j) {
                    const float f1j = __shfl_sync(FULL, cf1, j), f2j = __shfl_sync(FULL, cf2, j);
#pragma unroll
                    for (int q = 0; q < NS; ++q) {
                        const bool live = lane + 32 * q < nrows;
                        mybits[q] |= (live && av[q] == f1j) ? (1u << j) : 0u;
                        anyb = anyb || (live && bv[q] == f2j);
                    }
                }
                unsigned allbits = 0u;
                int nmatch = 0;
#pragma unroll
                for (int q = 0; q < NS; ++q) {
                    allbits |= mybits[q];
                    nmatch += __popc(mybits[q]);
                }
                const unsigned foundmask = __reduce_or_sync(FULL, allbits);
                const int total_matches = __reduce_add_sync(FULL, nmatch);
                const unsigned passmask = (np == 32) ? 0xffffffffu : ((1u << np) - 1u);
                const unsigned unfound = ~foundmask & passmask;
                const int nnew = __popc(unfound);
                if (__any_sync(FULL, anyb) || (total_matches != __popc(foundmask)) || (nrows + nnew > reg_cap) ||
                    (nrows + nnew > 32 * NS))
                    return false;
                // matched rows take their connection's update (:144-149); a row matches at most one connection
#pragma unroll
                for (int q = 0; q < NS; ++q) {
                    const int kk = mybits[q] ? (__ffs(mybits[q]) - 1) : 0;
                    const float g2 = __shfl_sync(FULL, cf2, kk), gadd = __shfl_sync(FULL, cadd, kk);
                    const bool wr = mybits[q] != 0u;
                    bv[q] = wr ? g2 : bv[q];
                    cn[q] = wr ? __fadd_rn(cn[q], 1.0f) : cn[q];
                    sc[q] = wr ? __fadd_rn(sc[q], gadd) : sc[q];
                }
                // unmatched connections append rows in connection order (:171-181): each posts its row at its
                // rank, the lane that owns that row index picks it up
                if (nnew) {
                    if ((unfound >> lane) & 1u) {
                        const int idx = nrows + __popc(unfound & ((1u << lane) - 1u));
                        s_stage[0][idx] = cf1;
                        s_stage[1][idx] = cf2;
                        s_stage[2][idx] = cnew;
                    }
                    __syncwarp();
#pragma unroll
                    for (int q = 0; q < NS; ++q) {
                        const int idx = 32 * q + lane;
                        const bool wr = idx >= nrows && idx < nrows + nnew;
                        av[q] = wr ? s_stage[0][idx] : av[q];
                        bv[q] = wr ? s_stage[1][idx] : bv[q];
                        sc[q] = wr ? s_stage[2][idx] : sc[q];
                        cn[q] = wr ? 2.0f : cn[q];
                    }
                    __syncwarp();
                    nrows += nnew;
                }
                return true;
            };
            while (limb < 17 && ws.k3_parallel_limbs && kpar < nconn) {
                const int np = min(32, nconn - kpar);
                bool ok;
                if (nrows + np <= 32) ok = limb_pass(std::integral_constant<int, 1>(), np);
                else if (nrows + np <= 64) ok = limb_pass(std::integral_constant<int, 2>(), np);
                else ok = limb_pass(std::integral_constant<int, S>(), np);
                if (!ok) break;                         // this pass (and the rest of the limb) one by one
                kpar += np;
            }
            for (int k = kpar; k < nconn; ++k) {
                bool bail;
                if (nrows <= 32) bail = step(std::integral_constant<int, 1>(), k);
                else if (nrows <= 64) bail = step(std::integral_constant<int, 2>(), k);
                else bail = step(std::integral_constant<int, S>(), k);
                if (bail) break;
            }
            // this limb's columns go back to the table (PA changes only where a row was created)
#pragma unroll
            for (int q = 0; q < S; ++q)
                if (lane + 32 * q < nrows) {
                    colA[lane + 32 * q] = av[q];
                    colB[lane + 32 * q] = bv[q];
                }
            __syncwarp();
        }
        float *col18 = rows + 18 * (size_t)MR, *col19 = rows + 19 * (size_t)MR;
#pragma unroll
        for (int q = 0; q < S; ++q)
            if (lane + 32 * q < nrows) {
                col18[lane + 32 * q] = sc[q];
                col19[lane + 32 * q] = cn[q];
            }
        __syncwarp();
    }
    for (int limb = limb0; limb < kLimbs; ++limb) {
        const int pa = c_limb_part[limb][0], pb = c_limb_part[limb][1];
        float *colA = rows + (size_t)pa * MR, *colB = rows + (size_t)pb * MR;
        float *col18 = rows + 18 * (size_t)MR, *col19 = rows + 19 * (size_t)MR;
        const int beg = s_lstart[limb], nconn = s_ccount[limb];
        for (int k = (limb == limb0) ? k0 : 0; k < nconn; ++k) {
            const unsigned long long rec = keys[beg + k];
            const int cid1 = (int)((rec >> 16) & 0xffffu), cid2 = (int)(rec & 0xffffu);
            const float cscore = __uint_as_float((uint32_t)(rec >> 32));
            const float f1 = (float)cid1, f2 = (float)cid2;
            int found = 0, r1 = 0, r2 = 0;
            for (int base = 0; base < nrows; base += 32) {
                const int r = base + lane;
                const bool hit = (r < nrows) && (colA[r] == f1 || colB[r] == f2);       // :137
                unsigned bits = __ballot_sync(0xffffffffu, hit);
                if (bits) {
                    if (found == 0) {
                        r1 = base + __ffs(bits) - 1;
                        unsigned rest = bits & (bits - 1);
                        if (rest) r2 = base + __ffs(rest) - 1;
                    } else if (found == 1) {
                        r2 = base + __ffs(bits) - 1;
                    }
                    found += __popc(bits);
                }
            }
            if (found == 1) {
                if (lane == 0 && colB[r1] != f2) {                                      // :144-149
                    colB[r1] = f2;
                    col19[r1] = __fadd_rn(col19[r1], 1.0f);
                    col18[r1] = __fadd_rn(col18[r1], __fadd_rn(pscore[cid2], cscore));
                }
            } else if (found == 2) {
                bool merge = false;
                if (th.mode == TFPD_MODE_UPSTREAM) {
                    // membership stays 0 iff no part k has both rows > 0 (:153-157)
                    bool both = (lane < kParts) && (rows[(size_t)lane * MR + r1] > 0.0f) &&
                                (rows[(size_t)lane * MR + r2] > 0.0f);
                    merge = (__ballot_sync(0xffffffffu, both) == 0u);
                }
                if (merge) {                                                           // :159-165
                    if (lane < kParts) {
                        float *cc = rows + (size_t)lane * MR;
                        cc[r1] = __fadd_rn(cc[r1], __fadd_rn(cc[r2], 1.0f));
                    } else if (lane == 18) {
                        col18[r1] = __fadd_rn(__fadd_rn(col18[r1], col18[r2]), cscore);
                    } else if (lane == 19) {
                        col19[r1] = __fadd_rn(col19[r1], col19[r2]);
                    }
                    __syncwarp();
                    // subset.erase(begin + r2): shift the later rows up by one, column by column
                    for (int c = 0; c < 20; ++c) {
                        float *cc = rows + (size_t)c * MR;
                        for (int base = r2; base < nrows - 1; base += 32) {
                            const int r = base + lane;
                            float v = 0.f;
                            if (r < nrows - 1) v = cc[r + 1];
                            __syncwarp();
                            if (r < nrows - 1) cc[r] = v;
                            __syncwarp();
                        }
                    }
                    --nrows;
                } else if (lane == 0) {                                                // :166-170
                    colB[r1] = f2;
                    col19[r1] = __fadd_rn(col19[r1], 1.0f);
                    col18[r1] = __fadd_rn(col18[r1], __fadd_rn(pscore[cid2], cscore));
                }
            } else if (found == 0 && limb < 17) {                                      // :171-181
                if (nrows < MR) {
                    if (lane < 20) {
                        float v = -1.0f;
                        if (lane == pa) v = f1;
                        if (lane == pb) v = f2;
                        if (lane == 19) v = 2.0f;
                        if (lane == 18) v = __fadd_rn(__fadd_rn(pscore[cid1], pscore[cid2]), cscore);
                        rows[(size_t)lane * MR + nrows] = v;
                    }
                    ++nrows;
                } else {
                    row_overflow = true;
                }
            } else if (found > 2) {
                if (lane == 0) atomicOr(&s_flags, TFPD_FLAG_FOUND_GT2);
            }
            __syncwarp();
        }
    }

    if (ws.dbg_times && tid == 0 && n < 64) ws.dbg_times[2048 + n * 8 + 5] = clock64();
    // ---- 4. filter (:185-189) + getter view ----
    const float *col18 = rows + 18 * (size_t)MR, *col19 = rows + 19 * (size_t)MR;
    const int16_t *pkx = s_pkx, *pky = s_pky;
    tfpd_human *out = ws.humans + (size_t)n * ws.max_humans;
    int nh = 0;
    bool human_overflow = false;
    for (int base = 0; base < nrows; base += 32) {
        const int r = base + lane;
        bool keep = false;
        if (r < nrows) {
            const float cntf = col19[r], sc = col18[r];
            keep = !(cntf < (float)th.part_cnt || __fdiv_rn(sc, cntf) < th.human_score);
        }
        const unsigned bits = __ballot_sync(0xffffffffu, keep);
        const int slot = nh + __popc(bits & ((1u << lane) - 1u));
        if (keep) {
            if (slot < ws.max_humans) {
                tfpd_human *hm = out + slot;
                hm->score = __fdiv_rn(col18[r], col19[r]);                             // get_score (:202-204)
#pragma unroll 6
                for (int p = 0; p < kParts; ++p) {
                    const int cid = (int)rows[(size_t)p * MR + r];                     // get_part_cid (:198-200)
                    hm->cid[p] = (int16_t)cid;
                    hm->x[p] = (cid >= 0) ? pkx[cid] : (int16_t)-1;
                    hm->y[p] = (cid >= 0) ? pky[cid] : (int16_t)-1;
                    hm->part_score[p] = (cid >= 0) ? pscore[cid] : 0.0f;
                }
            } else {
                human_overflow = true;
            }
        }
        nh += __popc(bits);
    }
    human_overflow = __any_sync(0xffffffffu, human_overflow);
    if (lane == 0) {
        int flags = s_flags | ws.counters[n].flags;
        if (row_overflow) flags |= TFPD_FLAG_ROW_OVERFLOW;
        if (human_overflow) flags |= TFPD_FLAG_HUMAN_OVERFLOW;
        ws.info[n].n_humans = min(nh, ws.max_humans);
        ws.info[n].n_cands = cnt;
        ws.info[n].n_rows = nrows;
        ws.info[n].flags = flags;
        if (ws.dbg_times && n < 64) {
            ws.dbg_times[2048 + n * 8 + 6] = clock64();
            int tc = 0;
            for (int l = 0; l < kLimbs; ++l) tc += s_ccount[l];
            ws.dbg_times[2048 + n * 8 + 7] = ((unsigned long long)tc << 32) | (unsigned)s_maxseg;   // connections | longest segment
        }
#ifdef TFPD_CHECKED
        TFPD_CHK(guard_ok(g_keys) && guard_ok(g_rows) && guard_ok(g_used) && guard_ok(g_end));
        TFPD_CHK(nrows >= 0 && nh >= 0 && cnt <= ws.max_cands);
#endif
    }
}

}  // namespace

size_t assemble_smem_bytes(const Workspace &ws) {
    return (size_t)ws.max_cands * 8 + (size_t)20 * ws.max_rows * 4 + (size_t)kLimbs * 2 * kBitmapWords * 4 +
           (size_t)ws.max_peaks * 8 + 4 * TFPD_GUARD_BYTES;
}

int launch_assemble(const Workspace &ws, int N, const Thresholds &th, cudaStream_t s) {
    size_t smem = assemble_smem_bytes(ws);
    static size_t attr_smem[64][2] = {};               // per device and variant: function attributes do not travel between devices
    int dev = 0;
    cudaGetDevice(&dev);
    // two CTAs per SM when two copies of the shared memory (dynamic + ~2 KB static + 1 KB the hardware reserves per
    // CTA) fit the SM's 228 KB; TFPD_K3_MINB=1 forces the one-per-SM build (A/B timing)
    static const int forced = getenv("TFPD_K3_MINB") ? atoi(getenv("TFPD_K3_MINB")) : 0;
    const bool two = forced ? forced >= 2 : 2 * (smem + 3 * 1024) <= (size_t)228 * 1024;
    const void *fn = two ? (const void *)k_assemble<2> : (const void *)k_assemble<1>;
    if (smem > 48 * 1024 && smem > attr_smem[dev & 63][two]) {
        if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return -1;
        cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        attr_smem[dev & 63][two] = smem;
    }
    if (two) k_assemble<2><<<N, kK3Threads, smem, s>>>(ws, th);
    else k_assemble<1><<<N, kK3Threads, smem, s>>>(ws, th);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace tfpd
