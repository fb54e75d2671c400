/*
 * oracle/paf_oracle.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of tf_pose/pafprocess/pafprocess.cpp (the swig-bound C++
 * decode): peak listing (:22-44), PAF line-integral candidate scoring
 * (:48-94, :216-237), per-limb sort + greedy matching (:96-124), subset
 * assembly (:127-183) and filtering (:185-189).  Re-entrant (no globals) and
 * with the two behaviours the reference binary can have made explicit:
 *
 *   mode 0 "asbuilt"  : `int membership;` at pafprocess.cpp:152 is never
 *                       initialised; GCC -O1..-O3 deletes the merge branch, so
 *                       found==2 always takes the else arm (:166-170).
 *   mode 1 "upstream" : membership starts at 0 (ildoonet upstream) -> merge.
 *
 * PINNED: tests/test_oracle.py runs this restatement against the
 * reference's own pafprocess.cpp compiled from /root/reference into
 * oracle/_ref/ (both modes) on seeded dense inputs, and against the committed
 * golden vectors in tests/golden/ that were produced by the reference's
 * PoseEstimator.estimate_paf (tests/golden/make_golden.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.
 */
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>

namespace {

constexpr int kParts = 18;      // pafprocess.h:11 NUM_PART
constexpr int kLimbs = 19;      // pafprocess.h:15 COCOPAIRS_SIZE
constexpr int kSteps = 10;      // pafprocess.h:13 STEP_PAF

// pafprocess.h:16-24 (limb -> PAF channel pair, limb -> part pair)
const int kLimbChan[kLimbs][2] = {{12, 13}, {20, 21}, {14, 15}, {16, 17}, {22, 23}, {24, 25}, {0, 1},
                                  {2, 3},   {4, 5},   {6, 7},   {8, 9},   {10, 11}, {28, 29}, {30, 31},
                                  {34, 35}, {32, 33}, {36, 37}, {18, 19}, {26, 27}};
const int kLimbPart[kLimbs][2] = {{1, 2},  {1, 5},   {2, 3},   {3, 4},  {5, 6},   {6, 7},  {1, 8},
                                  {8, 9},  {9, 10},  {1, 11},  {11, 12}, {12, 13}, {1, 0},  {0, 14},
                                  {14, 16}, {0, 15}, {15, 17}, {2, 16}, {5, 17}};

struct Params {
    float thresh_heat = 0.05f;          // pafprocess.h:6
    float thresh_vector_score = 0.05f;  // :7
    int thresh_vector_cnt1 = 6;         // :8
    int thresh_part_cnt = 4;            // :9
    float thresh_human_score = 0.3f;    // :10
};

struct PeakRec { int x, y; float score; int part; };
// Same 16-byte shape and comparator as the reference's ConnectionCandidate
// (pafprocess.h:38-43, pafprocess.cpp:239-241) so that std::sort performs the
// identical sequence of moves (it is unstable above 16 elements).
struct Cand { int i1, i2; float score; float etc; };
bool cand_desc(Cand a, Cand b) { return a.score > b.score; }
struct Conn { int cid1, cid2; float score; int i1, i2; };

struct Result {
    std::vector<PeakRec> peaks;                 // flat, id = index
    std::vector<Conn> conns[kLimbs];
    std::vector<Cand> cands[kLimbs];            // before sorting (a-major, b-minor)
    std::vector<std::vector<float>> rows;       // surviving subset rows (20 floats)
    std::vector<std::vector<float>> rows_all;   // before filtering
    int flags = 0;                              // bit0: exact score tie inside one limb's candidates
                                                // bit1: a connection met found>2 and was dropped
};

inline int round_half(float v) { return (int)((double)v + 0.5); }   // pafprocess.cpp:235-237

void run(Result &R, const float *peaks, const float *heat, const float *paf, int H, int W,
         int mode, int sort_mode, const Params &P)
{
    const int CP = 19, CH = 19, CF = 38;   // channel counts p3, h3, f3
    // ---- peak listing: part-major, then y, then x (:24-37) ----
    std::vector<int> first(kParts + 1, 0);
    for (int part = 0; part < kParts; ++part) {
        first[part] = (int)R.peaks.size();
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x)
                if (peaks[((long)y * W + x) * CP + part] > P.thresh_heat)
                    R.peaks.push_back({x, y, heat[((long)y * W + x) * CH + part], part});
    }
    first[kParts] = (int)R.peaks.size();

    // ---- per limb: candidates, sort, greedy (:48-124) ----
    for (int limb = 0; limb < kLimbs; ++limb) {
        const int pa = kLimbPart[limb][0], pb = kLimbPart[limb][1];
        const int na = first[pa + 1] - first[pa], nb = first[pb + 1] - first[pb];
        if (na == 0 || nb == 0) continue;
        const int chx = kLimbChan[limb][0], chy = kLimbChan[limb][1];
        std::vector<Cand> &cands = R.cands[limb];
        for (int ia = 0; ia < na; ++ia) {
            const PeakRec &A = R.peaks[first[pa] + ia];
            for (int ib = 0; ib < nb; ++ib) {
                const PeakRec &B = R.peaks[first[pb] + ib];
                float vx = (float)(B.x - A.x), vy = (float)(B.y - A.y);
                float norm = (float)std::sqrt(vx * vx + vy * vy);
                if (norm < 1e-12) continue;
                vx = vx / norm;
                vy = vy / norm;
                const float stepx = (B.x - A.x) / (float)kSteps;
                const float stepy = (B.y - A.y) / (float)kSteps;
                float total = 0.0f;
                int hits = 0;
                for (int i = 0; i < kSteps; ++i) {          // i = 0..9, end point excluded
                    int lx = round_half(A.x + i * stepx);
                    int ly = round_half(A.y + i * stepy);
                    const float *px = paf + ((long)ly * W + lx) * CF;
                    float s = vx * px[chx] + vy * px[chy];
                    total += s;
                    if (s > P.thresh_vector_score) ++hits;
                }
                // float/int -> float ; the min() term is all double (:83)
                float crit2 = total / kSteps + std::min(0.0, 0.5 * H / norm - 1.0);
                if (hits > P.thresh_vector_cnt1 && crit2 > 0)
                    cands.push_back({ia, ib, crit2, crit2 + A.score + B.score});
            }
        }
        std::vector<Cand> sorted = cands;
        // exact score ties: the reference's order is then decided by libstdc++'s
        // introsort internals -> flag the frame.
        {
            std::vector<float> sc;
            for (auto &c : sorted) sc.push_back(c.score);
            std::sort(sc.begin(), sc.end());
            for (size_t i = 1; i < sc.size(); ++i)
                if (sc[i] == sc[i - 1]) R.flags |= 1;
        }
        if (sort_mode == 0) {
            std::sort(sorted.begin(), sorted.end(), cand_desc);
        } else {   // the total order the CUDA path uses: score desc, idx1 asc, idx2 asc
            std::sort(sorted.begin(), sorted.end(), [](const Cand &a, const Cand &b) {
                if (a.score != b.score) return a.score > b.score;
                if (a.i1 != b.i1) return a.i1 < b.i1;
                return a.i2 < b.i2;
            });
        }
        std::vector<Conn> &conns = R.conns[limb];
        for (const Cand &c : sorted) {
            bool taken = false;
            for (const Conn &k : conns)
                if (k.i1 == c.i1 || k.i2 == c.i2) { taken = true; break; }
            if (taken) continue;
            conns.push_back({first[pa] + c.i1, first[pb] + c.i2, c.score, c.i1, c.i2});
        }
    }

    // ---- subset assembly (:127-183) ----
    std::vector<std::vector<float>> &S = R.rows_all;
    for (int limb = 0; limb < kLimbs; ++limb) {
        const int pa = kLimbPart[limb][0], pb = kLimbPart[limb][1];
        for (const Conn &cn : R.conns[limb]) {
            int found = 0, r1 = 0, r2 = 0;
            for (int r = 0; r < (int)S.size(); ++r)
                if (S[r][pa] == cn.cid1 || S[r][pb] == cn.cid2) {
                    if (found == 0) r1 = r;
                    if (found == 1) r2 = r;
                    ++found;
                }
            if (found > 2) R.flags |= 2;
            if (found == 1) {
                if (S[r1][pb] != cn.cid2) {
                    S[r1][pb] = cn.cid2;
                    S[r1][19] += 1;
                    S[r1][18] += R.peaks[cn.cid2].score + cn.score;
                }
            } else if (found == 2) {
                bool overlap = false;
                for (int k = 0; k < kParts; ++k)
                    if (S[r1][k] > 0 && S[r2][k] > 0) overlap = true;
                if (mode == 1 && !overlap) {
                    for (int k = 0; k < kParts; ++k) S[r1][k] += (S[r2][k] + 1);
                    S[r1][19] += S[r2][19];
                    S[r1][18] += S[r2][18];
                    S[r1][18] += cn.score;
                    S.erase(S.begin() + r2);
                } else {
                    S[r1][pb] = cn.cid2;
                    S[r1][19] += 1;
                    S[r1][18] += R.peaks[cn.cid2].score + cn.score;
                }
            } else if (found == 0 && limb < 17) {
                std::vector<float> row(20, -1.0f);
                row[pa] = cn.cid1;
                row[pb] = cn.cid2;
                row[19] = 2;
                row[18] = R.peaks[cn.cid1].score + R.peaks[cn.cid2].score + cn.score;
                S.push_back(row);
            }
        }
    }
    // ---- filter (:185-189); survivors keep creation order ----
    for (const auto &row : S)
        if (!(row[19] < P.thresh_part_cnt || row[18] / row[19] < P.thresh_human_score))
            R.rows.push_back(row);
}

}  // namespace

extern "C" {

void *paf_oracle_run(const float *peaks, const float *heat, const float *paf, int H, int W,
                     int mode, int sort_mode)
{
    Result *R = new Result();
    run(*R, peaks, heat, paf, H, W, mode, sort_mode, Params());
    return R;
}
void *paf_oracle_run_cfg(const float *peaks, const float *heat, const float *paf, int H, int W,
                         int mode, int sort_mode, float thresh_heat, float thresh_vec, int thresh_cnt1,
                         int thresh_part_cnt, float thresh_human)
{
    Params P;
    P.thresh_heat = thresh_heat; P.thresh_vector_score = thresh_vec; P.thresh_vector_cnt1 = thresh_cnt1;
    P.thresh_part_cnt = thresh_part_cnt; P.thresh_human_score = thresh_human;
    Result *R = new Result();
    run(*R, peaks, heat, paf, H, W, mode, sort_mode, P);
    return R;
}
void paf_oracle_free(void *h) { delete (Result *)h; }
int paf_oracle_flags(void *h) { return ((Result *)h)->flags; }
int paf_oracle_num_peaks(void *h) { return (int)((Result *)h)->peaks.size(); }
void paf_oracle_get_peaks(void *h, int *part, int *x, int *y, float *score)
{
    const Result *R = (Result *)h;
    for (size_t i = 0; i < R->peaks.size(); ++i) {
        part[i] = R->peaks[i].part; x[i] = R->peaks[i].x; y[i] = R->peaks[i].y; score[i] = R->peaks[i].score;
    }
}
int paf_oracle_num_cands(void *h, int limb) { return (int)((Result *)h)->cands[limb].size(); }
void paf_oracle_get_cands(void *h, int limb, int *i1, int *i2, float *score)
{
    const auto &v = ((Result *)h)->cands[limb];
    for (size_t i = 0; i < v.size(); ++i) { i1[i] = v[i].i1; i2[i] = v[i].i2; score[i] = v[i].score; }
}
int paf_oracle_num_conns(void *h, int limb) { return (int)((Result *)h)->conns[limb].size(); }
void paf_oracle_get_conns(void *h, int limb, int *cid1, int *cid2, float *score)
{
    const auto &v = ((Result *)h)->conns[limb];
    for (size_t i = 0; i < v.size(); ++i) { cid1[i] = v[i].cid1; cid2[i] = v[i].cid2; score[i] = v[i].score; }
}
int paf_oracle_num_humans(void *h) { return (int)((Result *)h)->rows.size(); }
void paf_oracle_get_human_rows(void *h, float *rows20)
{
    const auto &v = ((Result *)h)->rows;
    for (size_t i = 0; i < v.size(); ++i)
        for (int k = 0; k < 20; ++k) rows20[i * 20 + k] = v[i][k];
}
int paf_oracle_num_rows_unfiltered(void *h) { return (int)((Result *)h)->rows_all.size(); }

}  // extern "C"
