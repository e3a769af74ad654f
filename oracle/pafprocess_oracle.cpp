// ============================================================================================
// TEST INFRASTRUCTURE ONLY — CPU oracle for the keypoint-assignment path.
//
// A from-scratch restatement (flat arrays, one pass per stage) of the algorithm in the reference's
// utils/pafprocess/pafprocess.cpp.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load this library; the product (improved-body-parts_b200/) never does.
//
// Parity status: PINNED.  tests/test_oracle_vs_ref.py checks this file against the UNMODIFIED reference
// compiled into oracle/_ref/libpafprocess_ref.so (recipe: oracle/build_oracle.py) on the heatmap_test.npz
// fixture, on seeded synthetic crowds and on adversarial tie-heavy inputs, and tests/golden/*.npz freeze
// the reference's outputs so the same check runs where /root/reference does not exist.
//
// Build: g++ -O2 -std=c++14 -fPIC -shared -ffp-contract=off   (no -march: FMA contraction would change
// `a.x + i*STEP` — SURVEY.md §8(a) row A7).
//
// Every stage cites the reference lines it follows (paths relative to /root/reference/).
// ============================================================================================
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

namespace {

// ---- constants: utils/pafprocess/pafprocess.h:6-27 -------------------------------------------------
const float kLimbSampleThresh = 0.1f;      // THRESH_PAF_SCORE
const float kStepRatio = 0.8f;             // THRESH_PAF_STEP_RATIO
const int kMinParts = 2;                   // THRESH_PART_CNT
const float kMinHumanScore = 0.45f;        // THRESH_SKELETON_SCORE
const int kParts = 18;                     // NUM_PART
const int kMaxSteps = 20;                  // STEP_PAF
const int kLenRate = 16;                   // LIMB_LENGTH_RATE
const float kMergeTolerance = 0.7f;        // MIN_SCORE_TOLERANCE
const int kLimbs = 30;                     // COCOPAIRS_SIZE
const int kLimbEnds[kLimbs][2] = {         // COCOPAIRS == config/config.py:114-121 limb_from/limb_to
    {1, 0},  {1, 14}, {1, 15}, {1, 16}, {1, 17}, {0, 14},  {0, 15}, {14, 16}, {15, 17}, {1, 2},
    {2, 3},  {3, 4},  {1, 5},  {5, 6},  {6, 7},  {1, 8},   {8, 9},  {9, 10},  {1, 11},  {11, 12},
    {12, 13}, {0, 2}, {0, 5},  {2, 8},  {8, 12}, {5, 11},  {11, 9}, {16, 2},  {17, 5},  {8, 11}};

struct Cand {       // pafprocess.h:52-58 (ConnectionCandidate)
    int a, b;       // index inside the two per-part lists
    float limb;     // criterion2
    float rank;     // overall_score
    float len;
};
struct Link {       // pafprocess.h:60-67 (Connection)
    int a, b;
    float limb;
    int id_a, id_b;
    float len;
};
struct Body {       // the 20-slot ConnectedPeak row of pafprocess.h:36-45, unrolled into named fields
    int id[kParts];
    float sc[kParts];
    float total;    // slot 18 .score
    int count;      // slot 19 .count
    float reach;    // slot 19 .length
};
struct Pt {         // pafprocess.h:29-34
    int x, y;
    float score;
    int id;
};

// state that outlives a call, like the reference's two globals (pafprocess.cpp:16-17)
std::vector<Body> g_bodies;
std::vector<Pt> g_line;
std::vector<Link> g_links[kLimbs];
int g_ncand[kLimbs];
int g_dup[kLimbs];

inline int half_up(float v) { return (int)((double)v + 0.5); }  // pafprocess.cpp:329-331

// ---- OpenCV-native bicubic x`up` sample (cv2.resize INTER_CUBIC, A=-0.75, IPP off) -------------------------
// Used only by the "feature-resolution" entry point: value of pixel (X,Y) of cv2.resize(map, fx=up, fy=up)
// computed lazily.  Order of operations = OpenCV 4.x resize.cpp: HResizeCubic (left-assoc 4-tap sum per row)
// then VResizeCubicVec_32f (b0*S0 + (b1*S1 + (b2*S2 + b3*S3))).  Checked bit-for-bit against cv2 in
// tests/test_oracle_numpy.py.  Call sites replaced: evaluate.py:77-80, demo_image.py:94-97.
inline void cubic_w(float t, float *w) {
    const float A = -0.75f;
    w[0] = ((A * (t + 1) - 5 * A) * (t + 1) + 8 * A) * (t + 1) - 4 * A;
    w[1] = ((A + 2) * t - (A + 3)) * t * t + 1;
    w[2] = ((A + 2) * (1 - t) - (A + 3)) * (1 - t) * (1 - t) + 1;
    w[3] = 1.f - w[0] - w[1] - w[2];
}
inline void cubic_tap(int d, int up, int n, int *idx, float *w) {
    // fx = (dx + 0.5) * scale - 0.5 with scale = 1/up as double, then float (OpenCV resize.cpp, cubic branch)
    double scale = 1.0 / (double)up;
    float f = (float)((d + 0.5) * scale - 0.5);
    int s = (int)std::floor(f);
    f -= s;
    cubic_w(f, w);
    for (int k = 0; k < 4; k++) {
        int j = s - 1 + k;
        idx[k] = j < 0 ? 0 : (j >= n ? n - 1 : j);
    }
}
inline float cubic_sample(const float *plane, int h, int w, int up, int X, int Y) {
    int xi[4], yi[4];
    float xw[4], yw[4];
    cubic_tap(X, up, w, xi, xw);
    cubic_tap(Y, up, h, yi, yw);
    float r[4];
    for (int k = 0; k < 4; k++) {
        const float *row = plane + (size_t)yi[k] * w;
        r[k] = row[xi[0]] * xw[0] + row[xi[1]] * xw[1] + row[xi[2]] * xw[2] + row[xi[3]] * xw[3];
    }
    return r[0] * yw[0] + (r[1] * yw[1] + (r[2] * yw[2] + r[3] * yw[3]));
}

struct LimbSource {
    // mode 0: full-resolution HWC map (reference layout, pafprocess.cpp:9 PAF macro)
    // mode 1: feature-resolution planar (C,h,w) map sampled through cubic_sample
    int mode;
    const float *data;
    int H, W, C;  // mode 0: map dims; mode 1: feature dims h,w and channel count
    int up;
    float at(int limb, int x, int y) const {
        if (mode == 0) return data[limb + (size_t)C * (x + (size_t)W * y)];
        return cubic_sample(data + (size_t)limb * H * W, H, W, up, x, y);
    }
};

bool rank_ge(const Cand &p, const Cand &q) { return p.rank >= q.rank; }  // pafprocess.cpp:333-335 (non-strict!)

int run(int n_img, int n_row, int n_col, const float *peaks, const LimbSource &src, int min_img_size) {
    // ---- stage 1: bucket peaks by part, ids in arrival order: pafprocess.cpp:29-48 ----
    std::vector<Pt> per_part[kParts];
    int arrival = 0;
    for (int i = 0; i < n_img; i++)
        for (int j = 0; j < n_row; j++) {
            const float *r = peaks + (size_t)n_col * (j + (size_t)n_row * i);
            Pt p;
            p.id = arrival++;
            p.x = (int)r[0];
            p.y = (int)r[1];
            p.score = r[2];
            int part = (int)r[4];
            if (part < 0 || part >= kParts) return -2;  // reference: undefined behaviour
            per_part[part].push_back(p);
        }
    g_line.clear();
    for (int t = 0; t < kParts; t++) g_line.insert(g_line.end(), per_part[t].begin(), per_part[t].end());

    // ---- stage 2+3: score every pair, order, pick greedily: pafprocess.cpp:52-130 ----
    for (int limb = 0; limb < kLimbs; limb++) {
        g_links[limb].clear();
        g_ncand[limb] = 0;
        g_dup[limb] = 0;
        const std::vector<Pt> &A = per_part[kLimbEnds[limb][0]];
        const std::vector<Pt> &B = per_part[kLimbEnds[limb][1]];
        if (A.empty() && B.empty()) continue;  // :57
        std::vector<Cand> cands;
        for (int ia = 0; ia < (int)A.size(); ia++)
            for (int ib = 0; ib < (int)B.size(); ib++) {
                float dx = (float)(B[ib].x - A[ia].x);  // :68-69 int difference stored in float
                float dy = (float)(B[ib].y - A[ia].y);
                float len = (float)std::sqrt(dx * dx + dy * dy);  // :70
                if (len < 1e-12) continue;                         // :71
                int steps = std::min(half_up(len + 1), kMaxSteps);  // :73
                // get_paf_scores :311-327
                const float sx = (B[ib].x - A[ia].x) / float(steps - 1);
                const float sy = (B[ib].y - A[ia].y) / float(steps - 1);
                float sum = 0.0f;
                int above = 0;
                for (int s = 0; s < steps; s++) {
                    int px = half_up(A[ia].x + s * sx);
                    int py = half_up(A[ia].y + s * sy);
                    float v = src.at(limb, px, py);
                    sum += v;                          // :85 sequential float sum
                    if (v > kLimbSampleThresh) above++;  // :87
                }
                // :90 — the penalty term and the addition are evaluated in double, then narrowed
                float limb_score = sum / steps + std::min(0.0, 0.5 * min_img_size / len - 1.0);
                float need = steps * kStepRatio;  // :91
                if (above > need && limb_score > 0) {  // :93
                    Cand c;
                    c.a = ia;
                    c.b = ib;
                    c.limb = limb_score;
                    c.rank = 0.5f * limb_score + 0.25f * A[ia].score + 0.25f * B[ib].score;  // :94-96
                    c.len = len;
                    cands.push_back(c);
                }
            }
        g_ncand[limb] = (int)cands.size();
        {   // bookkeeping for the tests: does this limb have duplicate sort keys?
            std::vector<float> keys;
            for (size_t k = 0; k < cands.size(); k++) keys.push_back(cands[k].rank);
            std::sort(keys.begin(), keys.end());
            for (size_t k = 1; k < keys.size(); k++)
                if (keys[k] == keys[k - 1]) g_dup[limb] = 1;
        }
        std::sort(cands.begin(), cands.end(), rank_ge);  // :109 same libstdc++, same comparator
        int cap = (int)std::min(A.size(), B.size());     // :111
        std::vector<char> used_a(A.size(), 0), used_b(B.size(), 0);
        for (size_t k = 0; k < cands.size(); k++) {      // :113-129
            const Cand &c = cands[k];
            if (used_a[c.a] || used_b[c.b]) continue;
            used_a[c.a] = used_b[c.b] = 1;
            Link l;
            l.a = c.a;
            l.b = c.b;
            l.limb = c.limb;
            l.id_a = A[c.a].id;
            l.id_b = B[c.b].id;
            l.len = c.len;
            g_links[limb].push_back(l);
            if ((int)g_links[limb].size() >= cap) break;
        }
    }

    // ---- stage 4: grow / merge bodies: pafprocess.cpp:133-275 ----
    g_bodies.clear();
    for (int limb = 0; limb < kLimbs; limb++) {
        const int pa = kLimbEnds[limb][0], pb = kLimbEnds[limb][1];
        for (size_t li = 0; li < g_links[limb].size(); li++) {
            const Link &l = g_links[limb][li];
            int hits = 0, first = 0, second = 0;
            for (int b = 0; b < (int)g_bodies.size(); b++)  // :143-150
                if (g_bodies[b].id[pa] == l.id_a || g_bodies[b].id[pb] == l.id_b) {
                    if (hits == 0) first = b;
                    if (hits == 1) second = b;
                    hits++;
                }
            if (hits == 1) {  // :152-180
                Body &m = g_bodies[first];
                int gate = (int)(m.reach * kLenRate);  // :154 float product truncated
                if (m.id[pb] == -1 && gate > l.len) {
                    m.id[pb] = l.id_b;
                    m.sc[pb] = l.limb;
                    m.count += 1;
                    m.reach = std::max(m.reach, l.len);
                    m.total += g_line[l.id_b].score + l.limb;
                } else if (m.id[pb] != l.id_b && m.sc[pb] <= l.limb && gate > l.len) {
                    m.id[pb] = l.id_b;
                    m.sc[pb] = l.limb;
                    m.total -= g_line[m.id[pb]].score + m.sc[pb];  // :168 subtracts the NEW values
                    m.total += g_line[l.id_b].score + l.limb;
                    m.reach = std::max(m.reach, l.len);
                } else if (m.id[pb] == l.id_b && m.sc[pb] <= l.limb) {
                    m.id[pb] = l.id_b;
                    m.sc[pb] = l.limb;
                    m.total -= g_line[m.id[pb]].score + m.sc[pb];  // :177
                    m.total += g_line[l.id_b].score + l.limb;
                    m.reach = std::max(m.reach, l.len);
                }
            } else if (hits == 2) {  // :182-256
                Body &m = g_bodies[first];
                Body &o = g_bodies[second];
                int gate = (int)(m.reach * kLenRate);  // :186
                bool overlap = false;
                float low_m = 0.0f, low_o = 0.0f;
                for (int k = 0; k < kParts; k++) {  // :196-206 ("assigned" is id > 0)
                    bool hm = m.id[k] > 0, ho = o.id[k] > 0;
                    if (hm) low_m = (low_m == 0.0f) ? m.sc[k] : std::min(low_m, m.sc[k]);
                    if (ho) low_o = (low_o == 0.0f) ? o.sc[k] : std::min(low_o, o.sc[k]);
                    if (hm && ho) overlap = true;
                }
                if (!overlap) {  // :218-230
                    float floor_score = std::min(low_m, low_o) * kMergeTolerance;
                    if (l.limb >= floor_score || l.len < gate) {
                        for (int k = 0; k < kParts; k++) {
                            m.id[k] += (o.id[k] + 1);
                            m.sc[k] += (o.sc[k] + 1);
                        }
                        m.count += o.count;
                        m.reach = std::max(m.reach, l.len);
                        m.total += o.total + l.limb;
                        g_bodies.erase(g_bodies.begin() + second);
                    }
                }
                // overlap: DELETE_SHARED_JOINTS is const false (pafprocess.h:16) => nothing happens (:231-256)
            } else if (hits == 0) {  // :257-273
                Body nb;
                for (int k = 0; k < kParts; k++) {
                    nb.id[k] = -1;
                    nb.sc[k] = -1;
                }
                nb.id[pa] = l.id_a;
                nb.id[pb] = l.id_b;
                nb.sc[pa] = l.limb;
                nb.sc[pb] = l.limb;
                nb.count = 2;
                nb.reach = l.len;
                nb.total = g_line[l.id_a].score + g_line[l.id_b].score + l.limb;
                g_bodies.push_back(nb);
            }
            // hits >= 3: link silently dropped
        }
    }
    // ---- stage 5: filter: pafprocess.cpp:277-282 ----
    for (int b = (int)g_bodies.size() - 1; b >= 0; b--)
        if (g_bodies[b].count < kMinParts || g_bodies[b].total / g_bodies[b].count < kMinHumanScore)
            g_bodies.erase(g_bodies.begin() + b);
    return 0;
}

}  // namespace

extern "C" {

// pafprocess.h:70 — identical argument meaning
int oracle_process_paf(int p1, int p2, int p3, const float *peaks, int f1, int f2, int f3,
                       const float *pafmap, int min_img_size) {
    if (p3 < 5 || f3 < kLimbs) return -1;
    LimbSource s;
    s.mode = 0;
    s.data = pafmap;
    s.H = f1;
    s.W = f2;
    s.C = f3;
    s.up = 1;
    return run(p1, p2, p3, peaks, s, min_img_size);
}

// same algorithm, limb values read lazily from the feature-resolution planar map (30,h,w) through the
// OpenCV-order bicubic x`up` (equals indexing cv2.resize(pafs, fx=up, fy=up, INTER_CUBIC), evaluate.py:77-80)
int oracle_process_paf_feat(int p1, int p2, int p3, const float *peaks, int h, int w,
                            const float *limb_planes, int up, int min_img_size) {
    if (p3 < 5) return -1;
    LimbSource s;
    s.mode = 1;
    s.data = limb_planes;
    s.H = h;
    s.W = w;
    s.C = kLimbs;
    s.up = up;
    return run(p1, p2, p3, peaks, s, min_img_size);
}

int oracle_get_num_humans() { return (int)g_bodies.size(); }                                   // :287-289
int oracle_get_part_peak_id(int human, int part) { return g_bodies[human].id[part]; }          // :291-293
float oracle_get_score(int human) { return g_bodies[human].total / g_bodies[human].count; }    // :295-297
int oracle_get_part_x(int cid) { return g_line[cid].x; }                                       // :299-301
int oracle_get_part_y(int cid) { return g_line[cid].y; }                                       // :303-305
float oracle_get_part_score(int cid) { return g_line[cid].score; }                             // :307-309

// introspection for stage-level parity tests (no reference counterpart)
int oracle_num_candidates(int limb) { return g_ncand[limb]; }
int oracle_has_duplicate_keys(int limb) { return g_dup[limb]; }
int oracle_num_links(int limb) { return (int)g_links[limb].size(); }
void oracle_get_link(int limb, int k, int *id_a, int *id_b, float *score, float *len) {
    const Link &l = g_links[limb][k];
    *id_a = l.id_a;
    *id_b = l.id_b;
    *score = l.limb;
    *len = l.len;
}
void oracle_get_human_raw(int human, int *ids, float *scores, float *total, int *count, float *reach) {
    const Body &b = g_bodies[human];
    std::memcpy(ids, b.id, sizeof(b.id));
    std::memcpy(scores, b.sc, sizeof(b.sc));
    *total = b.total;
    *count = b.count;
    *reach = b.reach;
}
// bicubic sample exposed so the numpy oracle and the C oracle can be cross-checked
float oracle_cubic_sample(const float *plane, int h, int w, int up, int X, int Y) {
    return cubic_sample(plane, h, w, up, X, Y);
}
}
