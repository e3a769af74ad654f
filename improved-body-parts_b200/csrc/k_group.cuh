// K4b — person grouping: grow / merge partial skeletons limb by limb, then filter.  One warp per image.
//
// Reference: utils/pafprocess/pafprocess.cpp:133-282.  The chain is sequential by construction (every connection
// reads and mutates the table the next one scans, and a merge erases a row so later indices shift), so the
// parallelism is images across CTAs and, inside a step, lanes over the skeleton scan (32 rows per ballot), over the
// 18 part slots of a merge, and over the columns of the row shift that implements vector::erase.
//
// It runs as the tail of the limb kernel: the last of an image's 30 limb CTAs to finish (ticket counter) calls
// group_image() with the CTA's whole dynamic shared memory.  The connections of all 30 limbs are first copied into
// shared memory in one coalesced sweep (each carries its two peak scores, so the chain never touches global memory:
// a dependent global load per connection was what bounded the stand-alone kernel), then warp 0 walks the chain on the
// table ids[S][18], limb scores[S][18], total[S], count[S], reach[S].
#pragma once
#include "ibp_device.cuh"

namespace ibp {

struct GroupArgs {
    const DevLink *links;    // (batch, 30, cap)
    const int *link_count;   // (batch, 30)
    int cap;
    int max_bodies;
    int max_humans;
    int len_rate;
    float merge_tol;
    int min_parts;
    float min_score;
    ibp_human *humans;       // (batch, max_humans)
    int *n_humans;           // (batch)
    uint32_t *status;        // (batch)
};

constexpr size_t kBodyBytes = kParts * 8 + 12;  // shared bytes per row of the skeleton table

// All threads of the CTA call this; `smem` / `smem_bytes` is the CTA's dynamic shared memory (free for re-use).
__device__ __forceinline__ void group_image(const GroupArgs &a, int img, unsigned char *smem, size_t smem_bytes) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int S = a.max_bodies;
    int *ids = reinterpret_cast<int *>(smem);                         // [S][18]
    float *sc = reinterpret_cast<float *>(ids + (size_t)S * kParts);  // [S][18]
    float *total = sc + (size_t)S * kParts;                           // [S]
    int *count = reinterpret_cast<int *>(total + S);                  // [S]
    float *reach = reinterpret_cast<float *>(count + S);              // [S]
    size_t off = ((size_t)S * kBodyBytes + 15) & ~(size_t)15;
    int *loff = reinterpret_cast<int *>(smem + off);                  // [31] start of each limb's run in `slinks`
    off += 32 * sizeof(int);
    DevLink *slinks = reinterpret_cast<DevLink *>(smem + off);
    const int slots = (int)((smem_bytes - off) / sizeof(DevLink));

    // ---- stage the connections (written by other CTAs: read through L2) ----
    if (tid == 0) {
        int acc = 0;
        for (int l = 0; l < kLimbs; l++) {
            loff[l] = acc;
            acc += __ldcg(a.link_count + img * kLimbs + l);
        }
        loff[kLimbs] = acc;
    }
    __syncthreads();
    for (int l = warp; l < kLimbs; l += (int)(blockDim.x >> 5)) {
        const int n = loff[l + 1] - loff[l];
        const float2 *src = reinterpret_cast<const float2 *>(a.links + ((size_t)img * kLimbs + l) * a.cap);
        float2 *dst = reinterpret_cast<float2 *>(slinks + loff[l]);
        const int fit = max(0, min(n, slots - loff[l]));
        for (int i = lane; i < fit * 3; i += 32) dst[i] = __ldcg(src + i);  // 24-byte records as three float2
    }
    __syncthreads();
    if (warp != 0) return;

    int nb = 0;
    bool overflow = false;
    for (int limb = 0; limb < kLimbs; limb++) {
        const int pa = c_limb_a[limb], pb = c_limb_b[limb];
        const int base_slot = loff[limb];
        const int nl = loff[limb + 1] - base_slot;
        const DevLink *G = a.links + ((size_t)img * kLimbs + limb) * a.cap;
        auto fetch = [&](int li) -> DevLink {
            DevLink l;
            if (base_slot + li < slots) {
                l = slinks[base_slot + li];
            } else {  // more connections than the shared run holds: read the rest through L2
                const float2 *g2 = reinterpret_cast<const float2 *>(G + li);
                const float2 q0 = __ldcg(g2), q1 = __ldcg(g2 + 1), q2 = __ldcg(g2 + 2);
                l.id_a = __float_as_int(q0.x);
                l.id_b = __float_as_int(q0.y);
                l.limb = q1.x;
                l.len = q1.y;
                l.score_a = q2.x;
                l.score_b = q2.y;
            }
            return l;
        };
        DevLink nxt;
        if (nl > 0) nxt = fetch(0);
        for (int li = 0; li < nl; li++) {
            const DevLink l = nxt;
            if (li + 1 < nl) nxt = fetch(li + 1);  // off the dependent chain
            // :143-150 scan: first two matching rows, total number of matches.  Every lane also reads the fields of
            // its row that the one-match case needs, so that case continues from registers (one shuffle) instead of
            // a second round of dependent shared-memory loads.
            int hits = 0, first = 0, second = 0;
            int f_idb = -1;
            float f_scb = 0.f, f_reach = 0.f, f_total = 0.f;
            for (int base = 0; base < nb; base += 32) {
                const int s = base + lane;
                const bool valid = s < nb;
                const int ia = valid ? ids[s * kParts + pa] : -2;
                const int ibv = valid ? ids[s * kParts + pb] : -2;
                const float scb = valid ? sc[s * kParts + pb] : 0.f;
                const float rch = valid ? reach[s] : 0.f;
                const float tot = valid ? total[s] : 0.f;
                const uint32_t bal = __ballot_sync(0xffffffffu, valid && (ia == l.id_a || ibv == l.id_b));
                if (bal) {
                    const int p0 = __ffs(bal) - 1;
                    if (hits == 0) {
                        first = base + p0;
                        f_idb = __shfl_sync(0xffffffffu, ibv, p0);
                        f_scb = __shfl_sync(0xffffffffu, scb, p0);
                        f_reach = __shfl_sync(0xffffffffu, rch, p0);
                        f_total = __shfl_sync(0xffffffffu, tot, p0);
                        const uint32_t rest = bal & (bal - 1);
                        if (rest) second = base + __ffs(rest) - 1;
                    } else if (hits == 1) {
                        second = base + p0;
                    }
                    hits += __popc(bal);
                }
            }
            if (hits == 1) {  // :152-180
                if (lane == 0) {
                    const int f = first;
                    const float gate = __int2float_rn(__float2int_rz(__fmul_rn(f_reach, __int2float_rn(a.len_rate))));  // :154
                    const int cur_id = f_idb;
                    const float cur_sc = f_scb;
                    const float add = __fadd_rn(l.score_b, l.limb);
                    if (cur_id == -1 && gate > l.len) {
                        ids[f * kParts + pb] = l.id_b;
                        sc[f * kParts + pb] = l.limb;
                        count[f] += 1;
                        reach[f] = fmaxf(f_reach, l.len);
                        total[f] = __fadd_rn(f_total, add);
                    } else if ((cur_id != l.id_b && cur_sc <= l.limb && gate > l.len) || (cur_id == l.id_b && cur_sc <= l.limb)) {
                        ids[f * kParts + pb] = l.id_b;
                        sc[f * kParts + pb] = l.limb;
                        // :168-169 / :177-178 — subtracts the values just written, then adds them back
                        total[f] = __fadd_rn(__fsub_rn(f_total, add), add);
                        reach[f] = fmaxf(f_reach, l.len);
                    }
                }
                __syncwarp();
            } else if (hits == 2) {  // :182-256
                const int f = first, g = second;
                int idm = -1, ido = -1;
                float scm = 0.f, sco = 0.f;
                if (lane < kParts) {
                    idm = ids[f * kParts + lane];
                    ido = ids[g * kParts + lane];
                    scm = sc[f * kParts + lane];
                    sco = sc[g * kParts + lane];
                }
                const bool hm = idm > 0, ho = ido > 0;  // :198-199 peak id 0 counts as unassigned
                const bool overlap = __ballot_sync(0xffffffffu, hm && ho) != 0;
                if (!overlap) {
                    // :201-205 running minimum with the 0.0f sentinel, evaluated in part order on every lane
                    float low_m = 0.0f, low_o = 0.0f;
                    for (int k = 0; k < kParts; k++) {
                        const int im = __shfl_sync(0xffffffffu, idm, k), io = __shfl_sync(0xffffffffu, ido, k);
                        const float sm = __shfl_sync(0xffffffffu, scm, k), so = __shfl_sync(0xffffffffu, sco, k);
                        if (im > 0) low_m = (low_m == 0.0f) ? sm : (sm < low_m ? sm : low_m);
                        if (io > 0) low_o = (low_o == 0.0f) ? so : (so < low_o ? so : low_o);
                    }
                    const float floor_score = __fmul_rn((low_o < low_m ? low_o : low_m), a.merge_tol);  // :219
                    const float gate = __int2float_rn(__float2int_rz(__fmul_rn(reach[f], __int2float_rn(a.len_rate))));  // :186
                    if (l.limb >= floor_score || l.len < gate) {  // :220
                        if (lane < kParts) {  // :222-225
                            ids[f * kParts + lane] = idm + (ido + 1);
                            sc[f * kParts + lane] = __fadd_rn(scm, __fadd_rn(sco, 1.0f));
                        }
                        if (lane == 0) {  // :226-228
                            count[f] += count[g];
                            reach[f] = fmaxf(reach[f], l.len);
                            total[f] = __fadd_rn(total[f], __fadd_rn(total[g], l.limb));
                        }
                        __syncwarp();
                        // :229 erase row g: every lane shifts its own column down, rows in increasing order
                        for (int r = g; r < nb - 1; r++) {
                            if (lane < kParts) {
                                ids[r * kParts + lane] = ids[(r + 1) * kParts + lane];
                                sc[r * kParts + lane] = sc[(r + 1) * kParts + lane];
                            } else if (lane == kParts) {
                                total[r] = total[r + 1];
                            } else if (lane == kParts + 1) {
                                count[r] = count[r + 1];
                            } else if (lane == kParts + 2) {
                                reach[r] = reach[r + 1];
                            }
                        }
                        nb--;
                    }
                }
                // overlap: DELETE_SHARED_JOINTS is const false (pafprocess.h:16) => the branch at :231-256 does nothing
                __syncwarp();
            } else if (hits == 0) {  // :257-273
                if (nb < S) {
                    if (lane < kParts) {
                        ids[nb * kParts + lane] = (lane == pa) ? l.id_a : ((lane == pb) ? l.id_b : -1);
                        sc[nb * kParts + lane] = (lane == pa || lane == pb) ? l.limb : -1.0f;
                    }
                    if (lane == 0) {
                        count[nb] = 2;
                        reach[nb] = l.len;
                        total[nb] = __fadd_rn(__fadd_rn(l.score_a, l.score_b), l.limb);
                    }
                    nb++;
                } else {
                    overflow = true;
                }
                __syncwarp();
            }
            // hits >= 3: the connection is silently dropped
        }
    }

    // :277-282 filter (order preserved) + results of the getters (:287-297)
    int n_out = 0;
    ibp_human *out = a.humans + (size_t)img * a.max_humans;
    for (int base = 0; base < nb; base += 32) {
        const int s = base + lane;
        bool keep = false;
        if (s < nb) keep = !(count[s] < a.min_parts || __fdiv_rn(total[s], __int2float_rn(count[s])) < a.min_score);
        const uint32_t bal = __ballot_sync(0xffffffffu, keep);
        if (keep) {
            const int o = n_out + __popc(bal & ((1u << lane) - 1u));
            if (o < a.max_humans) {
                ibp_human hmn;
                for (int k = 0; k < kParts; k++) hmn.peak_id[k] = ids[s * kParts + k];
                hmn.score = __fdiv_rn(total[s], __int2float_rn(count[s]));
                hmn.total = total[s];
                hmn.count = count[s];
                hmn.reach = reach[s];
                out[o] = hmn;
            }
        }
        n_out += __popc(bal);
    }
    if (lane == 0) {
        a.n_humans[img] = min(n_out, a.max_humans);
        uint32_t st = 0;
        if (overflow) st |= IBP_ST_BODY_OVERFLOW;
        if (n_out > a.max_humans) st |= IBP_ST_HUMAN_OVERFLOW;
        if (st) atomicOr(a.status + img, st);
    }
}

}  // namespace ibp
