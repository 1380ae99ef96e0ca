// tb_trace_v8.cuh — k_trace_v6's rounds on the binary tree with the ray state that the rounds do not read moved out
// of the register file.
//
// Why. k_trace_v6 waits on its node fetch with 28 warps per SM (7 CTAs x 72 registers): issue-active 60 %, a third of
// all stall samples on the first use of the node load (profiles/r2_summary.md). Warps are what hides that latency,
// and registers are what limits the warps. Of the 72, ten hold state that only a triangle round (1.85 per ray) or the
// retirement of a ray reads: the offset origin, the shear constants and axis permutation of the watertight test, the
// closest hit's rank and tag, the slot. Here they live in shared memory, [word][thread] behind the stack rows
// (conflict-free: a warp reads 32 consecutive words), and "this lane owns a finished ray" is read off the stack
// pointer instead of the slot (a ray ends by popping the sentinel below entry 0: sp == bottom - ROW; an idle lane
// has sp == bottom).
//
// With the eighth CTA in (64 registers, no spills), three more changes against k_trace_v6, each measured on the B200
// with the hit records compared bit for bit (experiments/README.md, profiles/r2_summary.md section 6):
//   - a stack entry is the 4-byte child reference alone: half the shared memory per row (L1 keeps ~100 KB beside
//     8 CTAs' stacks), no entry distance — the ~5 % of entries that the closest hit has culled by the time they are
//     popped are fetched and fail both box tests;
//   - a round's bookkeeping (park, two ballots, triangle-round decision, exit test: 44 of ~95 warp instructions) is
//     paid once per TB_V8_STEPS node steps (step 6b);
//   - refills take queue positions from a block the warp reserved one refill earlier (tb_refill_position): no atomic
//     round trip on the refill's critical path.
// C3 1M: 4.95 (v6) -> 6.2 Grays/s, 0.51 of the probed L2 stream peak in algorithmic bytes.
//
// Same answer as every other trace kernel: same nodes, same tb_triangle, equal t -> smaller rank.
#pragma once

#ifndef TB_V8_CTAS_PER_SM
#define TB_V8_CTAS_PER_SM 8
#endif

#ifndef TB_V8_STEPS
#define TB_V8_STEPS 4           /* node steps per round, see step 6b: 1 / 2 / 3 / 4 = 5360 / 5849 / 5972 / 5977 Mrays/s on C3 1M (profiles/r3a, r3b);
                                   with a leaf parked at every step 4 steps give another 1 % there and on the C2 frame (profiles/r3f) */
#endif

// Pop: entry [sp - ROW] becomes the lane's node when the lane asked for one (cur == POP). Unconditional read (the row
// below entry 0 is the sentinel, the row below that is padding), resolved with selects.
template <bool SPILL, int ROW>
__device__ __forceinline__ void tb_v8_pop(int& cur, unsigned& sp, unsigned smem_end, const int* l_stack) {
    int e;
    if (SPILL && sp > smem_end) e = l_stack[(sp - smem_end) / ROW - 1];
    else e = tb_lds32(sp - ROW);
    const bool need = cur == TB_REF_POP;
    cur = need ? e : cur;
    sp = need ? sp - ROW : sp;
}

// Node step: both child boxes, near child next, far child pushed.
template <bool SPILL, int ROW>
__device__ __forceinline__ void tb_v8_node_step(const TbSelRay& sr, const uint4& qa, const uint4& qb, float cull_t, int& cur, unsigned& sp,
                                                unsigned smem_end, int* l_stack) {
    const int ref0 = (int)qb.z, ref1 = (int)qb.w;
    float tmin0, tmin1;
    bool hit0 = tb_box_sel(sr, qa.x, qa.y, qa.z, &tmin0);
    bool hit1 = tb_box_sel(sr, qa.w, qb.x, qb.y, &tmin1);
    hit0 = hit0 && !(tmin0 > cull_t);
    hit1 = hit1 && !(tmin1 > cull_t);
    const bool near1 = tmin1 < tmin0;
    const bool both = hit0 && hit1;
    const int far = near1 ? ref0 : ref1;
    if (SPILL && sp >= smem_end) l_stack[(sp - smem_end) / ROW] = far;
    else tb_sts32(sp, far);                                      // scratch row when !both
    sp += both ? ROW : 0;
    const int one = hit0 ? ref0 : (hit1 ? ref1 : TB_REF_POP);
    cur = both ? (near1 ? ref1 : ref0) : one;
}

template <bool COUNT, bool SPILL>
__global__ void __launch_bounds__(TB_TRACE_THREADS, TB_V8_CTAS_PER_SM) k_trace_v8(TbTraceArgs a) {
    extern __shared__ __align__(16) unsigned char s_v6_stack[];    // [stack_smem + 3][threads] x 4 B entries, [TB_LEAN_WORDS][threads] words, [warps] x 8 B refill state
    constexpr int ROW = TB_TRACE_THREADS * 4;                      // one row of stack entries: the child reference alone
    constexpr int LW = TB_TRACE_THREADS * 4;                       // one row of lean words
    int l_stack[SPILL ? TB_V6_LOCAL : 1];
    const int tid = threadIdx.x;
    const unsigned lane = tid & 31u;
    const unsigned FULL = 0xffffffffu;
    const int total = a.count_ptr ? *a.count_ptr : a.count;
    const unsigned bottom = (unsigned)__cvta_generic_to_shared(s_v6_stack) + 2 * ROW + tid * 4;
    tb_sts32(bottom - ROW, TB_REF_DONE);                           // sentinel below entry 0: an empty stack pops "finished"
    const unsigned smem_end = bottom + a.stack_smem * ROW;         // SPILL: entries from this address on live in local memory
    // lean words: 0-2 offset origin, 3-5 shear constants, 6 kx | ky << 2 | kz << 4, 7 best rank, 8 best tag, 9 slot
    const unsigned lean = (unsigned)__cvta_generic_to_shared(s_v6_stack) + (a.stack_smem + 3) * ROW + tid * 4;

    int   cur = TB_REF_DONE, pend = TB_REF_DONE;
    unsigned sp = bottom;                          // address of the next free entry; == bottom: the lane owns no ray
    float best_t = FLT_MAX;
    bool  drained = false;                         // the ray queue is empty: nothing to refill with, stay in the rounds to the end
    TbSelRay sr;
    memset(&sr, 0, sizeof sr);
    const unsigned wstate = (unsigned)__cvta_generic_to_shared(s_v6_stack) + (a.stack_smem + 3) * ROW + TB_LEAN_WORDS * LW + (tid >> 5) * 8;
    int pre = tb_refill_init(lane, wstate, a.cursor);             // lane 0: first queue position of the block reserved ahead
    unsigned n_nodes = 0, n_tris = 0;
    unsigned n_rounds = 0, n_culled = 0, n_second = 0, n_parked = 0, n_empty = 0, n_tri_rounds = 0, n_refills = 0, n_refilled = 0;

    while (true) {
        // ---- refill idle lanes (warp converged here)
        const bool idle = cur == TB_REF_DONE && pend == TB_REF_DONE;
        const unsigned idle_mask = __ballot_sync(FULL, idle);
        if (idle_mask && !drained) {
            const int want = __popc(idle_mask);
            if (COUNT) ++n_refills;
            const int q = tb_refill_position(idle_mask, lane, wstate, pre, a.cursor, total, drained);
            if (idle) {
                if (q < total) {
                    const int slot = a.queue ? a.queue[q] : q;
                    const float4 o4 = tb_ld_stream(a.ray_o + slot), d4 = tb_ld_stream(a.ray_d + slot);
                    const V3 d = v3(d4.x, d4.y, d4.z);
                    const TbRay r = tb_ray_setup(v3(o4.x, o4.y, o4.z), d);
                    sr = tb_sel_ray(r, d, a.grid_lo, a.grid_cell);
                    tb_sts32(lean + 0 * LW, __float_as_int(r.o.x)); tb_sts32(lean + 1 * LW, __float_as_int(r.o.y)); tb_sts32(lean + 2 * LW, __float_as_int(r.o.z));
                    tb_sts32(lean + 3 * LW, __float_as_int(r.sx));  tb_sts32(lean + 4 * LW, __float_as_int(r.sy));  tb_sts32(lean + 5 * LW, __float_as_int(r.sz));
                    tb_sts32(lean + 6 * LW, r.kx | (r.ky << 2) | (r.kz << 4));
                    tb_sts32(lean + 7 * LW, -1);
                    tb_sts32(lean + 9 * LW, slot);
                    cur = 0; sp = bottom;
                    best_t = FLT_MAX;
                    if (COUNT) ++n_refilled;
                }
            }
        }
        if (__ballot_sync(FULL, cur != TB_REF_DONE || pend != TB_REF_DONE) == 0u) break;

        // ---- rounds
        while (true) {
            // entries and boxes beyond this distance cannot hold the closest hit (inf while there is none: culls nothing);
            // the margin keeps candidates with an EQUAL t alive for the smaller-rank rule
            const float cull_t = best_t + best_t * 1e-5f;
            // 1. park a leaf (a further leaf waits in `cur` until the parking slot is free)
            {
                const bool is_leaf = (unsigned)cur > (unsigned)TB_REF_POP;       // ~rank: 0x80000002 .. 0xffffffff
                const bool park = is_leaf && pend == TB_REF_DONE;
                pend = park ? cur : pend;
                cur = park ? TB_REF_POP : cur;
            }
            // 2. pop. An entry carries no distance: one that the closest hit has culled since it was pushed is fetched
            //    and fails both box tests (1.2 of 23 node fetches per ray on C3 1M) — the price of the half-size stack.
            tb_v8_pop<SPILL, ROW>(cur, sp, smem_end, l_stack);
            // 3. what does this round do?
            const bool is_node = cur >= 0;
            const unsigned node_mask = __ballot_sync(FULL, is_node);
            const unsigned pend_mask = __ballot_sync(FULL, pend != TB_REF_DONE);
            const bool tri_round = pend_mask != 0u && (__popc(pend_mask) >= a.tri_min || node_mask == 0u);
            const bool do_tri = tri_round && pend != TB_REF_DONE;
            if (COUNT) {
                ++n_rounds;
                if (tri_round) ++n_tri_rounds;
                if (!is_node) {
                    if (cur == TB_REF_POP) ++n_culled;
                    else if (cur != TB_REF_DONE) ++n_second;
                    else if (pend != TB_REF_DONE) ++n_parked;
                    else ++n_empty;
                }
            }
            // 4. all loads of the round first, so node and triangle latencies overlap
            float4 ta, tb, tc;
            uint4 qa, qb;
            if (is_node) tb_ld_qnode(a.qnodes + (size_t)cur * 2, qa, qb);
            if (do_tri) {
                const float4* tp = a.tris + (size_t)(~pend) * 3;
                ta = __ldg(tp); tb = __ldg(tp + 1); tc = __ldg(tp + 2);
            }
            // 5. node step
            if (is_node) {
                if (COUNT) ++n_nodes;
                tb_v8_node_step<SPILL, ROW>(sr, qa, qb, cull_t, cur, sp, smem_end, l_stack);
            }
            // 6. triangle step for the parked leaves: the watertight test's ray constants come back from shared memory
            if (do_tri) {
                const int rank = ~pend;
                if (COUNT) ++n_tris;
                TbRay r;
                r.o = v3(__int_as_float(tb_lds32(lean + 0 * LW)), __int_as_float(tb_lds32(lean + 1 * LW)), __int_as_float(tb_lds32(lean + 2 * LW)));
                r.inv = v3s(0.f);
                r.sx = __int_as_float(tb_lds32(lean + 3 * LW)); r.sy = __int_as_float(tb_lds32(lean + 4 * LW)); r.sz = __int_as_float(tb_lds32(lean + 5 * LW));
                const int kpack = tb_lds32(lean + 6 * LW);
                r.kx = kpack & 3; r.ky = (kpack >> 2) & 3; r.kz = (kpack >> 4) & 3;
                float t, u, v;
                if (tb_triangle(r, v3(ta.x, ta.y, ta.z), v3(tb.x, tb.y, tb.z), v3(tc.x, tc.y, tc.z), &t, &u, &v)) {
                    bool better = t < best_t;
                    if (t == best_t) {                                           // exact tie: the smaller rank wins
                        const int best_rank = tb_lds32(lean + 7 * LW);
                        better = best_rank >= 0 && rank < best_rank;
                    }
                    if (better) {
                        best_t = t;
                        tb_sts32(lean + 7 * LW, rank); tb_sts32(lean + 8 * LW, __float_as_int(tb.w));
                        const int slot = tb_lds32(lean + 9 * LW);
                        tb_st_stream(a.hit_out + slot, make_float4(t, a.emit_prim ? ta.w : __int_as_float(rank), a.emit_prim ? u : tb.w, v));
                    }
                }
                pend = TB_REF_DONE;
            }
            // 6b. further node steps of the round: a round's bookkeeping (park, two ballots, the triangle-round decision, the
            //     exit test: 44 of its ~95 warp instructions, ncu r2y) is paid once for TB_V8_STEPS node steps. A lane whose
            //     step ended on a leaf parks it if the parking slot is free and pops; with the slot taken it sits the rest
            //     of the round out (triangle rounds are decided once per round).
#pragma unroll
            for (int step = 1; step < TB_V8_STEPS; ++step) {
                const float cull2 = best_t + best_t * 1e-5f;                 // the triangle step may have moved the closest hit
                {                                                            // park, as at the top of the round
                    const bool is_leaf = (unsigned)cur > (unsigned)TB_REF_POP;
                    const bool park = is_leaf && pend == TB_REF_DONE;
                    pend = park ? cur : pend;
                    cur = park ? TB_REF_POP : cur;
                }
                tb_v8_pop<SPILL, ROW>(cur, sp, smem_end, l_stack);
                const bool is_node2 = cur >= 0;
                if (COUNT) {
                    ++n_rounds;
                    if (!is_node2) {
                        if (cur != TB_REF_DONE) ++n_second;
                        else if (pend != TB_REF_DONE) ++n_parked;
                        else ++n_empty;
                    }
                }
                if (is_node2) {
                    uint4 qc, qd;
                    tb_ld_qnode(a.qnodes + (size_t)cur * 2, qc, qd);
                    if (COUNT) ++n_nodes;
                    tb_v8_node_step<SPILL, ROW>(sr, qc, qd, cull2, cur, sp, smem_end, l_stack);
                }
            }
            // 7. leave the rounds when the warp is done or has too few rays left to be worth a round
            const unsigned live = __ballot_sync(FULL, cur != TB_REF_DONE || pend != TB_REF_DONE);
            if (live == 0u || (!drained && __popc(live) < a.refetch_below)) break;
        }

        // ---- retire finished rays (warp converged)
        const bool finished = cur == TB_REF_DONE && pend == TB_REF_DONE && sp != bottom;
        if (__any_sync(FULL, finished)) {
            int preset = -1, slot = -1;
            if (finished) {
                slot = tb_lds32(lean + 9 * LW);
                if (best_t < FLT_MAX) preset = (tb_lds32(lean + 8 * LW) >> 8) & 0xff;
                else tb_st_stream(a.hit_out + slot, make_float4(-1.f, __int_as_float(-1), 0.f, 0.f));
                sp = bottom;
            }
            if (a.matq) {
#pragma unroll
                for (int p = 0; p < TB_PRESET_COUNT; ++p) {
                    const unsigned m = __ballot_sync(FULL, finished && preset == p);
                    if (m) {
                        const int leader = __ffs(m) - 1;
                        int base = 0;
                        if ((int)lane == leader) base = atomicAdd(a.matq_count + p, __popc(m));
                        base = __shfl_sync(FULL, base, leader);
                        if (finished && preset == p) a.matq[(size_t)p * a.matq_stride + base + __popc(m & ((1u << lane) - 1u))] = slot;
                    }
                }
            }
        }
    }
    if (COUNT) {
        atomicAdd(a.visit_counters, (unsigned long long)n_nodes);
        atomicAdd(a.visit_counters + 1, (unsigned long long)n_tris);
        if (a.round_stats) {
            const unsigned v[10] = { n_rounds, n_nodes, n_culled, n_second, n_parked, n_empty, n_tri_rounds, n_tris, n_refills, n_refilled };
#pragma unroll
            for (int k = 0; k < 10; ++k) atomicAdd(a.round_stats + k, (unsigned long long)v[k]);
        }
    }
}
