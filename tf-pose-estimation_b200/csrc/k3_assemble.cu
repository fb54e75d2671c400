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

#include "tfpd_internal.cuh"

namespace tfpd {

namespace {

// COCOPAIRS (pafprocess.h:16-19) as compile-time functions: the register-resident assembly below needs the
// part columns of a limb as constants so that the row registers are indexed statically.
__host__ __device__ constexpr int limb_pa(int l) {
    constexpr int t[kLimbs] = {1, 1, 2, 3, 5, 6, 1, 8, 9, 1, 11, 12, 1, 0, 14, 0, 15, 2, 5};
    return t[l];
}
__host__ __device__ constexpr int limb_pb(int l) {
    constexpr int t[kLimbs] = {2, 5, 3, 4, 6, 7, 8, 9, 10, 11, 12, 13, 0, 14, 16, 15, 17, 16, 17};
    return t[l];
}
constexpr int kRegRowsMax = 64;             // subset rows held in registers (2 per lane) before spilling to smem

constexpr int kK3Threads = 640;             // 20 warps: 19 greedy walkers + sort helpers
constexpr int kBitmapWords = kMaxPeaksHard / 32;   // 128

__global__ void __launch_bounds__(kK3Threads) k_assemble(Workspace ws, Thresholds th) {
    extern __shared__ __align__(16) unsigned char sm_raw[];
    unsigned long long *keys = reinterpret_cast<unsigned long long *>(sm_raw);            // [max_cands]
    float *rows = reinterpret_cast<float *>(keys + ws.max_cands);                        // [20][max_rows] SoA
    uint32_t *used = reinterpret_cast<uint32_t *>(rows + 20 * (size_t)ws.max_rows);      // [19][2][128]
    // the frame's peak table, staged once: assembly and the getter view index it by cid (a dependent global
    // load per connection was the latency chain of this kernel)
    float *s_pscore = reinterpret_cast<float *>(used + kLimbs * 2 * kBitmapWords);       // [max_peaks]
    int16_t *s_pkx = reinterpret_cast<int16_t *>(s_pscore + ws.max_peaks);               // [max_peaks]
    int16_t *s_pky = s_pkx + ws.max_peaks;                                               // [max_peaks]
    __shared__ int s_lstart[kLimbs + 1];
    __shared__ int s_ccount[kLimbs];
    __shared__ int s_first[kHeatC + 1];
    __shared__ int s_flags;

    const int n = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int MR = ws.max_rows;
    const int raw = ws.counters[n].n_cand;
    const int cnt = min(raw, ws.max_cands);
    int np2 = 32;
    while (np2 < cnt) np2 <<= 1;
    for (int i = tid; i < np2; i += kK3Threads)
        keys[i] = (i < cnt) ? ws.cand_key[(size_t)n * ws.max_cands + i] : ~0ull;
    for (int i = tid; i < kLimbs * 2 * kBitmapWords; i += kK3Threads) used[i] = 0u;
    if (tid <= kHeatC) s_first[tid] = ws.part_first[n * 20 + tid];
    if (tid <= kLimbs) s_lstart[tid] = cnt;
    if (tid < kLimbs) s_ccount[tid] = 0;
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

    // ---- 1. sort ----
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
    // limb segments + exact-tie detection
    for (int i = tid; i < cnt; i += kK3Threads) {
        int limb = (int)(keys[i] >> 56);
        int prev = (i == 0) ? -1 : (int)(keys[i - 1] >> 56);
        for (int q = prev + 1; q <= limb; ++q) s_lstart[q] = i;
        if (i > 0 && (keys[i] >> 24) == (keys[i - 1] >> 24)) atomicOr(&s_flags, TFPD_FLAG_SCORE_TIE);
    }
    __syncthreads();

    // ---- 2. greedy matching, one warp per limb (lane 0 walks; the walk is inherently serial) ----
    int16_t *c1 = ws.conn_cid1 + (size_t)n * ws.max_cands;
    int16_t *c2 = ws.conn_cid2 + (size_t)n * ws.max_cands;
    float *cs = ws.conn_score + (size_t)n * ws.max_cands;
    if (warp < kLimbs && lane == 0) {
        const int limb = warp;
        const int beg = s_lstart[limb], end = s_lstart[limb + 1];
        uint32_t *ua = used + (size_t)limb * 2 * kBitmapWords, *ub = ua + kBitmapWords;
        const int fa = s_first[c_limb_part[limb][0]], fb = s_first[c_limb_part[limb][1]];
        int nconn = 0;
        for (int i = beg; i < end; ++i) {
            const unsigned long long key = keys[i];
            const int a = (int)((key >> 12) & 0xfff), b = (int)(key & 0xfff);
            if ((ua[a >> 5] >> (a & 31)) & 1u) continue;        // peak_id1 already assigned (:104-108)
            if ((ub[b >> 5] >> (b & 31)) & 1u) continue;        // peak_id2 already assigned (:110-114)
            ua[a >> 5] |= 1u << (a & 31);
            ub[b >> 5] |= 1u << (b & 31);
            c1[beg + nconn] = (int16_t)(fa + a);                // conn.cid1 = peak_a_list[idx1].id (:122)
            c2[beg + nconn] = (int16_t)(fb + b);
            const uint32_t sbits = 0xffffffffu - (uint32_t)((key >> 24) & 0xffffffffu);
            cs[beg + nconn] = __uint_as_float(sbits);
            // the same record for phase 3, packed over a key this walk has already consumed (beg+nconn <= i)
            keys[beg + nconn] = ((unsigned long long)sbits << 32) | ((unsigned long long)(uint32_t)(fa + a) << 16) |
                                (unsigned long long)(uint32_t)(fb + b);
            ++nconn;
        }
        s_ccount[limb] = nconn;
    }
    __syncthreads();
    if (tid < kLimbs) {
        ws.conn_start[n * 20 + tid] = s_lstart[tid];
        ws.conn_count[n * 20 + tid] = s_ccount[tid];
    }
    if (warp != 0) return;

    // ---- 3. subset assembly (warp 0) ----
    // Fast path for the first kRegRowsMax rows: row r belongs to lane r & 31, slot r >> 5.  A limb only ever
    // looks at two columns of a row (its parts PA, PB) and updates PB, the score and the count, so per limb each
    // lane pulls its rows' PA / PB entries out of the smem table ONCE, keeps score and count in registers for the
    // whole assembly, and a connection is then two compares per lane, a ballot and an update inside the owning
    // lane -- no shared-memory round trip on the serial chain (the all-smem loop below spends ~900 cycles per
    // connection on them) and, unlike a per-limb instantiation with statically indexed row registers, one small
    // loop body that stays in the instruction cache.  When a 65th row is needed, the registers go back to the smem
    // rows and the generic loop below continues from the same (limb, connection).
    const float *pscore = s_pscore;
    const unsigned FULL = 0xffffffffu;
    int nrows = 0;
    bool row_overflow = false;
    int limb0 = kLimbs, k0 = 0;                       // where the generic loop resumes (kLimbs = nowhere)
    {
        const int reg_cap = min(min(ws.k3_reg_rows, kRegRowsMax), MR);
        // a row that does not exist yet reads as "no part anywhere": creating one (:172-180) is then 4 writes
        for (int c = 0; c < kParts; ++c) {
            if (lane < MR) rows[(size_t)c * MR + lane] = -1.0f;
            if (lane + 32 < MR) rows[(size_t)c * MR + lane + 32] = -1.0f;
        }
        __syncwarp();
        float sc0 = 0.f, sc1 = 0.f, cn0 = 0.f, cn1 = 0.f;            // columns 18 / 19 of this lane's two rows
        const bool own0 = lane < MR, own1 = lane + 32 < MR;
        for (int limb = 0; limb < kLimbs && limb0 == kLimbs; ++limb) {
            const int pa = c_limb_part[limb][0], pb = c_limb_part[limb][1];
            float *colA = rows + (size_t)pa * MR, *colB = rows + (size_t)pb * MR;
            float a0 = own0 ? colA[lane] : -1.0f, a1 = own1 ? colA[lane + 32] : -1.0f;
            float b0v = own0 ? colB[lane] : -1.0f, b1v = own1 ? colB[lane + 32] : -1.0f;
            const int beg = s_lstart[limb], nconn = s_ccount[limb];
            for (int k = 0; k < nconn; ++k) {
                const unsigned long long rec = keys[beg + k];
                const int cid1 = (int)((rec >> 16) & 0xffffu), cid2 = (int)(rec & 0xffffu);
                const float cscore = __uint_as_float((uint32_t)(rec >> 32));
                const float f1 = (float)cid1, f2 = (float)cid2;
                const float ps1 = pscore[cid1], ps2 = pscore[cid2];
                const unsigned m0 = __ballot_sync(FULL, (lane < nrows) && (a0 == f1 || b0v == f2));     // :137
                unsigned m1 = 0u;
                if (nrows > 32) m1 = __ballot_sync(FULL, (lane + 32 < nrows) && (a1 == f1 || b1v == f2));
                const int found = __popc(m0) + __popc(m1);
                if (found == 0) {
                    if (limb < 17) {                                                            // :171-181
                        if (nrows >= reg_cap) {
                            if (nrows >= MR) {
                                row_overflow = true;
                                continue;
                            }
                            limb0 = limb;               // continue in shared memory from this connection
                            k0 = k;
                            break;
                        }
                        const float sc = __fadd_rn(__fadd_rn(ps1, ps2), cscore);
                        if (lane == (nrows & 31)) {
                            if (nrows < 32) {
                                a0 = f1;
                                b0v = f2;
                                sc0 = sc;
                                cn0 = 2.0f;
                            } else {
                                a1 = f1;
                                b1v = f2;
                                sc1 = sc;
                                cn1 = 2.0f;
                            }
                        }
                        ++nrows;
                    }
                } else if (found <= 2) {
                    const int r1 = m0 ? (__ffs(m0) - 1) : (32 + __ffs(m1) - 1);
                    bool merged = false;
                    if (found == 2 && th.mode == TFPD_MODE_UPSTREAM) {
                        // :151-165, rare: done on the smem rows (one copy of the code), registers re-read after
                        const unsigned rest0 = m0 & (m0 - 1u);
                        const int r2 = (m0 == 0u) ? (32 + __ffs(m1 & (m1 - 1u)) - 1)
                                                  : (rest0 ? (__ffs(rest0) - 1) : (32 + __ffs(m1) - 1));
                        float *col18 = rows + 18 * (size_t)MR, *col19 = rows + 19 * (size_t)MR;
                        if (lane < nrows) {
                            colA[lane] = a0;
                            colB[lane] = b0v;
                            col18[lane] = sc0;
                            col19[lane] = cn0;
                        }
                        if (lane + 32 < nrows) {
                            colA[lane + 32] = a1;
                            colB[lane + 32] = b1v;
                            col18[lane + 32] = sc1;
                            col19[lane + 32] = cn1;
                        }
                        __syncwarp();
                        const bool both = (lane < kParts) && (rows[(size_t)lane * MR + r1] > 0.0f) &&
                                          (rows[(size_t)lane * MR + r2] > 0.0f);
                        merged = (__ballot_sync(FULL, both) == 0u);                    // :153-157
                        if (merged) {
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
                                const int ra = lane, rb = lane + 32;
                                float v0 = 0.f, v1 = 0.f;
                                if (ra >= r2 && ra < nrows) v0 = (ra + 1 < nrows) ? cc[ra + 1] : -1.0f;
                                if (rb >= r2 && rb < nrows) v1 = (rb + 1 < nrows) ? cc[rb + 1] : -1.0f;
                                __syncwarp();
                                if (ra >= r2 && ra < nrows) cc[ra] = v0;
                                if (rb >= r2 && rb < nrows) cc[rb] = v1;
                            }
                            --nrows;
                            __syncwarp();
                            a0 = own0 ? colA[lane] : -1.0f;
                            a1 = own1 ? colA[lane + 32] : -1.0f;
                            b0v = own0 ? colB[lane] : -1.0f;
                            b1v = own1 ? colB[lane + 32] : -1.0f;
                            sc0 = own0 ? col18[lane] : 0.f;
                            sc1 = own1 ? col18[lane + 32] : 0.f;
                            cn0 = own0 ? col19[lane] : 0.f;
                            cn1 = own1 ? col19[lane + 32] : 0.f;
                        }
                    }
                    if (!merged && lane == (r1 & 31)) {                           // :144-149 / :166-170
                        if (r1 < 32) {
                            if (found == 2 || b0v != f2) {
                                b0v = f2;
                                cn0 = __fadd_rn(cn0, 1.0f);
                                sc0 = __fadd_rn(sc0, __fadd_rn(ps2, cscore));
                            }
                        } else {
                            if (found == 2 || b1v != f2) {
                                b1v = f2;
                                cn1 = __fadd_rn(cn1, 1.0f);
                                sc1 = __fadd_rn(sc1, __fadd_rn(ps2, cscore));
                            }
                        }
                    }
                } else {
                    if (lane == 0) atomicOr(&s_flags, TFPD_FLAG_FOUND_GT2);
                }
            }
            // this limb's columns go back to the table (PA changes only where a row was created)
            if (lane < nrows) {
                colA[lane] = a0;
                colB[lane] = b0v;
            }
            if (lane + 32 < nrows) {
                colA[lane + 32] = a1;
                colB[lane + 32] = b1v;
            }
            __syncwarp();
        }
        float *col18 = rows + 18 * (size_t)MR, *col19 = rows + 19 * (size_t)MR;
        if (lane < nrows) {
            col18[lane] = sc0;
            col19[lane] = cn0;
        }
        if (lane + 32 < nrows) {
            col18[lane + 32] = sc1;
            col19[lane + 32] = cn1;
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
    }
}

}  // namespace

size_t assemble_smem_bytes(const Workspace &ws) {
    return (size_t)ws.max_cands * 8 + (size_t)20 * ws.max_rows * 4 + (size_t)kLimbs * 2 * kBitmapWords * 4 +
           (size_t)ws.max_peaks * 8;
}

int launch_assemble(const Workspace &ws, int N, const Thresholds &th, cudaStream_t s) {
    size_t smem = assemble_smem_bytes(ws);
    static size_t attr_smem = 0;
    if (smem > 48 * 1024 && smem > attr_smem) {
        if (cudaFuncSetAttribute(k_assemble, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return -1;
        attr_smem = smem;
    }
    k_assemble<<<N, kK3Threads, smem, s>>>(ws, th);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace tfpd
