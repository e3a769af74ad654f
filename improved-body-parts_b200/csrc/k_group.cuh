// K4b — person grouping: grow / merge partial skeletons limb by limb, then filter.  One warp per image.
//
// Reference: utils/pafprocess/pafprocess.cpp:133-282.  The chain is sequential by construction (every connection
// reads and mutates the table the next one scans), so the parallelism is images across CTAs and, inside a step,
// lanes over the skeleton scan (32 rows per ballot) and over the 18 part slots of a merge.  What is left is a
// latency chain of one warp, so the code is organised around the number of dependent steps per connection:
//
//   * It runs as the tail of the limb kernel: the last of an image's 30 limb CTAs to finish (ticket counter) calls
//     group_image() with the CTA's whole dynamic shared memory.  The connections of all 30 limbs are first copied into
//     shared memory in one coalesced sweep; each carries its two peak scores and its length gate, so the chain never
//     touches global memory (a dependent global load per connection bounded the first version of this step).
//   * Row layout: slot[S][18] = {peak id, limb score} pairs (one 64-bit load gives both), row[S] = {total, gate,
//     reach, count} (one 128-bit load).  Every lane loads the three values of ITS row during the scan; the lane that
//     owns the single matching row then updates it from its own registers — no second round of loads, no shuffles.
//   * The reference's `(int)(length * 16) > conn.length` test (:154, :186) needs a float->int->float round trip on
//     the chain.  x -> (float)(int)(x*16) is monotonic and a row's length is the running maximum of connection
//     lengths, so gate(row) = max over its connections of gate(connection): the limb kernel computes the latter once
//     per connection and the row keeps the running maximum — the same bits, no conversion on the chain.
//   * vector::erase (:229) shifts every later row down, which only matters through the ORDER of the rows (the scan
//     keeps the first two matches, the output keeps row order).  Rows here are never moved: an erased row gets ids
//     that match nothing and count -1 (so the final filter drops it); surviving rows keep their relative order.
//     max_bodies therefore bounds the number of rows ever created, not the number alive at one time.
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

// Slots of one row are 19 (not 18) int2 apart: with an odd stride in 8-byte units the 32 lanes of a warp, each reading
// the same part of a different row, hit different banks.
constexpr int kSlotStride = kParts + 1;
constexpr size_t kBodyBytes = kSlotStride * 8 + 16;  // shared bytes per row of the skeleton table
constexpr int kDeadId = -3;
constexpr int kGroupParallelMin = 4;  // connections of one limb type from which the one-per-lane path pays                      // id in every slot of an erased row: equals no peak id, not -1

// pafprocess.cpp:154 / :186  min_len = (int)(length * LIMB_LENGTH_RATE), later compared with a float
__device__ __forceinline__ float length_gate(float len, int len_rate) {
    return __int2float_rn(__float2int_rz(__fmul_rn(len, __int2float_rn(len_rate))));
}

// All threads of the CTA call this; `smem` / `smem_bytes` is the CTA's dynamic shared memory (free for re-use).
__device__ __forceinline__ void group_image(const GroupArgs &a, int img, unsigned char *smem, size_t smem_bytes,
                                            unsigned long long *tl = nullptr) {  // tl: experiments, two %globaltimer stamps
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int S = a.max_bodies;
    float4 *row = reinterpret_cast<float4 *>(smem);                       // [S] {total, gate, reach, count bits}
    int2 *slot = reinterpret_cast<int2 *>(smem + (size_t)S * 16);         // [S][18] {id, limb score bits}
    size_t off = (size_t)S * kBodyBytes;
    int *loff = reinterpret_cast<int *>(smem + off);                      // [31] start of each limb's run in `slinks`
    off += 32 * sizeof(int);
    DevLink *slinks = reinterpret_cast<DevLink *>(smem + off);
    const int slots = (int)((smem_bytes - off) / sizeof(DevLink));

    // ---- stage the connections (written by other CTAs: read through L2) ----
    if (tid < 32) {
        const int n = (tid < kLimbs) ? __ldcg(a.link_count + img * kLimbs + tid) : 0;
        int incl = n;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        loff[tid] = incl - n;  // loff[30] = total, loff[31] unused
    }
    // the first 32 rows can be scanned without a bounds test: give them ids that match nothing
    for (int i = tid; i < 32 * kSlotStride; i += (int)blockDim.x) slot[i] = make_int2(kDeadId, 0);
    __syncthreads();
    for (int l = warp; l < kLimbs; l += (int)(blockDim.x >> 5)) {
        const int n = loff[l + 1] - loff[l];
        const float4 *src = reinterpret_cast<const float4 *>(a.links + ((size_t)img * kLimbs + l) * a.cap);
        float4 *dst = reinterpret_cast<float4 *>(slinks + loff[l]);
        const int fit = max(0, min(n, slots - loff[l]));
        for (int i = lane; i < fit * 2; i += 32) dst[i] = __ldcg(src + i);  // 32-byte records as two float4
    }
    __syncthreads();
    if (warp != 0) return;
    if (tl && lane == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl[0]));

    // :182-256 two rows matched: merge row g into row f when the rule says so.  The whole warp, lane k = part k.
    auto merge_rows = [&](int f, int g, float link_score, float link_len, float link_gate) {
        int2 vm = make_int2(-1, 0), vo = make_int2(-1, 0);
        if (lane < kParts) {
            vm = slot[f * kSlotStride + lane];
            vo = slot[g * kSlotStride + lane];
        }
        const float4 rf = row[f], rg = row[g];
        const int idm = vm.x, ido = vo.x;
        const float scm = __int_as_float(vm.y), sco = __int_as_float(vo.y);
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
            if (link_score >= floor_score || link_len < rf.y) {  // :220, rf.y = (int)(length*16) of row f (:186)
                if (lane < kParts) {
                    // :222-225 ids and scores are ADDED (+1 undoes the -1 of an empty slot)
                    slot[f * kSlotStride + lane] = make_int2(idm + (ido + 1), __float_as_int(__fadd_rn(scm, __fadd_rn(sco, 1.0f))));
                    slot[g * kSlotStride + lane] = make_int2(kDeadId, 0);  // :229 erase
                }
                if (lane == 0) {  // :226-228
                    float4 nr;
                    nr.x = __fadd_rn(rf.x, __fadd_rn(rg.x, link_score));
                    nr.y = fmaxf(rf.y, link_gate);
                    nr.z = fmaxf(rf.z, link_len);
                    nr.w = __int_as_float(__float_as_int(rf.w) + __float_as_int(rg.w));
                    row[f] = nr;
                    row[g] = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
                }
            }
        }
        // overlap: DELETE_SHARED_JOINTS is const false (pafprocess.h:16) => the branch at :231-256 does nothing
        __syncwarp();
    };

    int nb = 0;
    bool overflow = false;
    const int my_loff = loff[lane];
    for (int limb = 0; limb < kLimbs; limb++) {
        const int pa = c_limb_a[limb], pb = c_limb_b[limb];
        const int base_slot = __shfl_sync(0xffffffffu, my_loff, limb);  // lane l holds loff[l]: no shared-memory round trip
        const int nl = __shfl_sync(0xffffffffu, my_loff, limb + 1) - base_slot;
        const DevLink *G = a.links + ((size_t)img * kLimbs + limb) * a.cap;
        auto fetch = [&](int li) -> DevLink {
            union {
                float4 q[2];
                DevLink l;
            } u;
            if (base_slot + li < slots) {
                const float4 *p = reinterpret_cast<const float4 *>(slinks + base_slot + li);
                u.q[0] = p[0];
                u.q[1] = p[1];
            } else {  // more connections than the shared run holds: read the rest through L2
                const float4 *p = reinterpret_cast<const float4 *>(G + li);
                u.q[0] = __ldcg(p);
                u.q[1] = __ldcg(p + 1);
            }
            return u.l;
        };
        // ---- up to 32 rows: lane r keeps row r's two slots of this limb and its totals in registers ----
        // The dependent chain of the common cases (extend a row, open a row) then never goes through shared memory,
        // whose latency under the other CTA's staging traffic is what a connection otherwise costs.  Shared memory
        // is brought up to date at the end of the limb (and around a merge, which touches all 18 slots of two rows).
        bool limb_done = false;
        if (nl > 0 && nb + nl <= 32 && base_slot + nl <= slots) {
            limb_done = true;
            int my_a = slot[lane * kSlotStride + pa].x;  // rows >= nb hold ids that match nothing
            int2 my_b = slot[lane * kSlotStride + pb];
            float4 my_row = row[lane];
            const float4 *lp = reinterpret_cast<const float4 *>(slinks + base_slot);
            union LinkBits {
                float4 q[2];
                DevLink l;
            };
            LinkBits nx;
            nx.q[0] = lp[0];
            nx.q[1] = lp[1];
            for (int li = 0; li < nl; li++) {
                const DevLink l = nx.l;
                if (li + 1 < nl) {
                    nx.q[0] = lp[2 * (li + 1)];
                    nx.q[1] = lp[2 * (li + 1) + 1];
                }
                const uint32_t bal = __ballot_sync(0xffffffffu, my_a == l.id_a || my_b.x == l.id_b);  // :143-150
                if (bal == 0) {  // :257-273 a new row, owned by lane nb
                    if (lane < kParts) {
                        const bool end = (lane == pa || lane == pb);
                        slot[nb * kSlotStride + lane] = make_int2((lane == pa) ? l.id_a : ((lane == pb) ? l.id_b : -1),
                                                                  __float_as_int(end ? l.limb : -1.0f));
                    }
                    if (lane == nb) {
                        my_a = l.id_a;
                        my_b = make_int2(l.id_b, __float_as_int(l.limb));
                        my_row = make_float4(__fadd_rn(__fadd_rn(l.score_a, l.score_b), l.limb), l.gate, l.len, __int_as_float(2));
                    }
                    nb++;
                } else if ((bal & (bal - 1)) == 0) {  // :152-180 exactly one row, on the lane that owns it
                    if (lane == __ffs(bal) - 1) {
                        const float gate = my_row.y;  // :154
                        const int cur_id = my_b.x;
                        const float cur_sc = __int_as_float(my_b.y);
                        const float add = __fadd_rn(l.score_b, l.limb);
                        const bool grow = (cur_id == -1) & (gate > l.len);
                        const bool swap = !grow & (cur_sc <= l.limb) & ((cur_id == l.id_b) | ((cur_id != l.id_b) & (gate > l.len)));
                        if (grow | swap) {
                            my_b = make_int2(l.id_b, __float_as_int(l.limb));
                            // swap (:168-169 / :177-178) subtracts the values just written, then adds them back
                            my_row.x = grow ? __fadd_rn(my_row.x, add) : __fadd_rn(__fsub_rn(my_row.x, add), add);
                            my_row.y = fmaxf(gate, l.gate);
                            my_row.z = fmaxf(my_row.z, l.len);
                            my_row.w = __int_as_float(__float_as_int(my_row.w) + (grow ? 1 : 0));
                        }
                    }
                } else if (__popc(bal) == 2) {  // :182-256 through shared memory, then back into the registers
                    slot[lane * kSlotStride + pb] = my_b;
                    row[lane] = my_row;
                    __syncwarp();
                    const int f = __ffs(bal) - 1;
                    merge_rows(f, __ffs(bal & (bal - 1)) - 1, l.limb, l.len, l.gate);
                    my_a = slot[lane * kSlotStride + pa].x;
                    my_b = slot[lane * kSlotStride + pb];
                    my_row = row[lane];
                }
                // three or more rows: the connection is silently dropped
            }
            slot[lane * kSlotStride + pb] = my_b;
            row[lane] = my_row;
            __syncwarp();
        }
        DevLink nxt;
        if (nl > 0 && !limb_done) nxt = fetch(0);
        for (int li = 0; li < nl && !limb_done; li++) {
            const DevLink l = nxt;
            if (li + 1 < nl) nxt = fetch(li + 1);  // off the dependent chain
            // :143-150 scan: first two matching rows, total number of matches
            int hits = 0, first = 0, second = 0;
            int2 my_pb;
            float4 my_row;
            if (nb <= 32) {
                // one row per lane; rows >= nb still hold the never-matching initial ids, so no bounds test
                const int ia = slot[lane * kSlotStride + pa].x;
                my_pb = slot[lane * kSlotStride + pb];
                my_row = row[lane];
                const uint32_t bal = __ballot_sync(0xffffffffu, ia == l.id_a || my_pb.x == l.id_b);
                hits = __popc(bal);
                first = __ffs(bal) - 1;
                second = __ffs(bal & (bal - 1)) - 1;
            } else {
                my_pb = make_int2(-1, 0);
                my_row = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int base = 0; base < nb; base += 32) {
                    const int s = base + lane;
                    int ia = kDeadId;
                    int2 pbv = make_int2(kDeadId, 0);
                    float4 rw = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (s < nb) {
                        ia = slot[s * kSlotStride + pa].x;
                        pbv = slot[s * kSlotStride + pb];
                        rw = row[s];
                    }
                    const uint32_t bal = __ballot_sync(0xffffffffu, ia == l.id_a || pbv.x == l.id_b);
                    if (bal) {
                        const int p0 = __ffs(bal) - 1;
                        if (hits == 0) {
                            first = base + p0;
                            my_pb = pbv;  // the owner lane of row `first` keeps that row's values
                            my_row = rw;
                            const uint32_t rest = bal & (bal - 1);
                            if (rest) second = base + __ffs(rest) - 1;
                        } else if (hits == 1) {
                            second = base + p0;
                        }
                        hits += __popc(bal);
                    }
                }
            }
            if (hits == 1) {  // :152-180, on the lane that owns the row, from the values it loaded during the scan
                const int f = first;
                const float gate = my_row.y;  // :154
                const int cur_id = my_pb.x;
                const float cur_sc = __int_as_float(my_pb.y);
                const float add = __fadd_rn(l.score_b, l.limb);
                const bool grow = (cur_id == -1) & (gate > l.len);
                const bool swap = !grow & (cur_sc <= l.limb) & ((cur_id == l.id_b) | ((cur_id != l.id_b) & (gate > l.len)));
                if ((lane == (f & 31)) & (grow | swap)) {
                    slot[f * kSlotStride + pb] = make_int2(l.id_b, __float_as_int(l.limb));
                    float4 nr;
                    // swap (:168-169 / :177-178) subtracts the values just written, then adds them back
                    nr.x = grow ? __fadd_rn(my_row.x, add) : __fadd_rn(__fsub_rn(my_row.x, add), add);
                    nr.y = fmaxf(gate, l.gate);
                    nr.z = fmaxf(my_row.z, l.len);
                    nr.w = __int_as_float(__float_as_int(my_row.w) + (grow ? 1 : 0));
                    row[f] = nr;
                }
                __syncwarp();
            } else if (hits == 2) {  // :182-256
                merge_rows(first, second, l.limb, l.len, l.gate);
            } else if (hits == 0) {  // :257-273
                if (nb < S) {
                    if (lane < kParts) {
                        const bool end = (lane == pa || lane == pb);
                        slot[nb * kSlotStride + lane] = make_int2((lane == pa) ? l.id_a : ((lane == pb) ? l.id_b : -1),
                                                             __float_as_int(end ? l.limb : -1.0f));
                    }
                    if (lane == 0)
                        row[nb] = make_float4(__fadd_rn(__fadd_rn(l.score_a, l.score_b), l.limb), l.gate, l.len, __int_as_float(2));
                    nb++;
                } else {
                    overflow = true;
                }
                __syncwarp();
            }
            // hits >= 3: the connection is silently dropped
        }
    }

    if (tl && lane == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl[1]));
    // :277-282 filter (order preserved; erased rows carry count -1) + results of the getters (:287-297)
    int n_out = 0;
    ibp_human *out = a.humans + (size_t)img * a.max_humans;
    for (int base = 0; base < nb; base += 32) {
        const int s = base + lane;
        bool keep = false;
        float4 rw = make_float4(0.f, 0.f, 0.f, 0.f);
        int cnt = -1;
        if (s < nb) {
            rw = row[s];
            cnt = __float_as_int(rw.w);
            keep = cnt >= 0 && !(cnt < a.min_parts || __fdiv_rn(rw.x, __int2float_rn(cnt)) < a.min_score);
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, keep);
        if (keep) {
            const int o = n_out + __popc(bal & ((1u << lane) - 1u));
            if (o < a.max_humans) {
                ibp_human hmn;
                for (int k = 0; k < kParts; k++) hmn.peak_id[k] = slot[s * kSlotStride + k].x;
                hmn.score = __fdiv_rn(rw.x, __int2float_rn(cnt));
                hmn.total = rw.x;
                hmn.count = cnt;
                hmn.reach = rw.z;
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
