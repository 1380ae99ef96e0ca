// tb_trace_v3.cuh — closest-hit kernel, third structure: warp-synchronous ROUNDS with postponed leaves.
//
// Why (ncu n1, source page of the while-while kernel on incoherent rays): the node loop ran with 7.1 of
// 32 lanes per instruction and 43 % of all stall samples sat on the first use of the node loads — the
// kernel is bound by memory LATENCY at low memory-level parallelism, because lanes that reached a leaf
// idle until the slowest lane of the warp reaches one. (Neither the instruction count nor the width
// of the loads mattered: FMA boxes and 256-bit loads changed nothing on those rays.)
//
// Here every lane that owns a ray issues one memory request per round:
//   - a lane at an inner node fetches it, tests the two child boxes and descends / pushes / pops;
//   - a lane that arrives at a leaf parks it in `pend` and keeps walking (its cull distance is one
//     leaf stale, which can only cost extra nodes, never a wrong answer);
//   - the watertight triangle test (~150 instructions) runs for all parked leaves at once, when at
//     least TB_TRI_MIN lanes have one or no lane can do node work; its loads are issued in the same
//     round as the node loads, so the two latencies overlap.
// One pop per round, no inner loops: a popped entry that is already culled just costs that lane a round.
#pragma once

#define TB_REF_POP  ((int)0x80000001)             /* lane must pop its stack next round */

template <bool COUNT, bool EXACT, bool QUANT>
__global__ void __launch_bounds__(TB_TRACE_THREADS, TB_TRACE_CTAS_PER_SM) k_trace_v3(TbTraceArgs a) {
    __shared__ int   s_ref[TB_STACK_SMEM][TB_TRACE_THREADS];
    __shared__ float s_tmin[TB_STACK_SMEM][TB_TRACE_THREADS];
    int   l_ref[TB_STACK_LOCAL];
    float l_tmin[TB_STACK_LOCAL];
    const int tid = threadIdx.x;
    const unsigned lane = tid & 31u;
    const unsigned FULL = 0xffffffffu;
    const int total = a.count_ptr ? *a.count_ptr : a.count;

    int   slot = -1, cur = TB_REF_DONE, pend = TB_REF_DONE, sp = 0;
    int   best_rank = -1;
    float best_t = FLT_MAX, cull_t = FLT_MAX, best_u = 0.f, best_v = 0.f;
    TbRay r;
    r.o = r.inv = v3s(0.f); r.sx = r.sy = r.sz = 0.f; r.kx = r.ky = r.kz = 0;
    TbFastRay fr;
    fr.ix = fr.iy = fr.iz = fr.ox = fr.oy = fr.oz = 0.f;
    bool exhausted = false;
    unsigned n_nodes = 0, n_tris = 0;

    while (true) {
        // ---- refill idle lanes (warp converged here)
        const bool idle = cur == TB_REF_DONE && pend == TB_REF_DONE;
        const unsigned idle_mask = __ballot_sync(FULL, idle);
        if (idle_mask && !exhausted) {
            const int want = __popc(idle_mask);
            const int leader = __ffs(idle_mask) - 1;
            int base = 0;
            if ((int)lane == leader) base = atomicAdd(a.cursor, want);
            base = __shfl_sync(FULL, base, leader);
            if (base + want >= total) exhausted = true;
            if (idle) {
                const int q = base + __popc(idle_mask & ((1u << lane) - 1u));
                if (q < total) {
                    slot = a.queue ? a.queue[q] : q;
                    const float4 o4 = a.ray_o[slot], d4 = a.ray_d[slot];
                    const V3 d = v3(d4.x, d4.y, d4.z);
                    r = tb_ray_setup(v3(o4.x, o4.y, o4.z), d);
                    if (QUANT) fr = tb_quant_ray(r, d, a.grid_lo, a.grid_cell);
                    else if (!EXACT) fr = tb_fast_ray(r, d);
                    cur = 0; sp = 0;
                    best_rank = -1; best_t = FLT_MAX; cull_t = FLT_MAX;
                }
            }
        }
        if (__ballot_sync(FULL, cur != TB_REF_DONE || pend != TB_REF_DONE) == 0u) break;

        // ---- rounds
        while (true) {
            // 1. one pop for lanes that need it (culled entries cost the lane this round)
            if (cur == TB_REF_POP) {
                if (sp == 0) cur = TB_REF_DONE;
                else {
                    --sp;
                    int n_; float tm_;
                    if (sp < TB_STACK_SMEM) { n_ = s_ref[sp][tid]; tm_ = s_tmin[sp][tid]; }
                    else { n_ = l_ref[sp - TB_STACK_SMEM]; tm_ = l_tmin[sp - TB_STACK_SMEM]; }
                    cur = tm_ > cull_t ? TB_REF_POP : n_;
                }
            }
            // 2. park a leaf and keep walking (a second leaf waits in `cur` until the parked one is tested)
            if (cur < 0 && cur != TB_REF_DONE && cur != TB_REF_POP && pend == TB_REF_DONE) { pend = cur; cur = TB_REF_POP; }
            // 3. what does this round do?
            const bool is_node = cur >= 0;
            const unsigned node_mask = __ballot_sync(FULL, is_node);
            const unsigned pend_mask = __ballot_sync(FULL, pend != TB_REF_DONE);
            const bool tri_round = pend_mask != 0u && (__popc(pend_mask) >= a.tri_min || node_mask == 0u);
            const bool do_tri = tri_round && pend != TB_REF_DONE;
            // 4. all loads of the round first, so node and triangle latencies overlap
            float4 q0, q1, q2, q3, ta, tb, tc;
            uint4 qa, qb;
            if (is_node) {
                if (QUANT) {
                    const uint4* np = a.qnodes + (size_t)cur * 2;
                    qa = __ldg(np); qb = __ldg(np + 1);
                } else {
                    const float4* np = a.nodes + (size_t)cur * 4;
                    q0 = __ldg(np); q1 = __ldg(np + 1); q2 = __ldg(np + 2); q3 = __ldg(np + 3);
                }
            }
            if (do_tri) {
                const float4* tp = a.tris + (size_t)(~pend) * 3;
                ta = __ldg(tp); tb = __ldg(tp + 1); tc = __ldg(tp + 2);
            }
            // 5. node step
            if (is_node) {
                if (COUNT) ++n_nodes;
                int ref0, ref1;
                float tmin0, tmin1;
                bool hit0, hit1;
                if (QUANT) {
                    ref0 = (int)qb.z; ref1 = (int)qb.w;
                    hit0 = tb_box_fast(fr, tb_qlo(qa.x), tb_qhi(qa.x), tb_qlo(qa.y), tb_qhi(qa.y), tb_qlo(qa.z), tb_qhi(qa.z), &tmin0);
                    hit1 = tb_box_fast(fr, tb_qlo(qa.w), tb_qhi(qa.w), tb_qlo(qb.x), tb_qhi(qb.x), tb_qlo(qb.y), tb_qhi(qb.y), &tmin1);
                } else if (EXACT) {
                    ref0 = __float_as_int(q3.x); ref1 = __float_as_int(q3.y);
                    hit0 = tb_box(r, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, &tmin0);
                    hit1 = tb_box(r, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, &tmin1);
                } else {
                    ref0 = __float_as_int(q3.x); ref1 = __float_as_int(q3.y);
                    hit0 = tb_box_fast(fr, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, &tmin0);
                    hit1 = tb_box_fast(fr, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, &tmin1);
                }
                hit0 = hit0 && !(tmin0 > cull_t);
                hit1 = hit1 && !(tmin1 > cull_t);
                const bool near1 = tmin1 < tmin0;
                if (hit0 && hit1) {
                    const int   far_ref = near1 ? ref0 : ref1;
                    const float far_tmin = near1 ? tmin0 : tmin1;
                    if (sp < TB_STACK_SMEM) { s_ref[sp][tid] = far_ref; s_tmin[sp][tid] = far_tmin; }
                    else { l_ref[sp - TB_STACK_SMEM] = far_ref; l_tmin[sp - TB_STACK_SMEM] = far_tmin; }
                    ++sp;
                }
                cur = (hit0 && hit1) ? (near1 ? ref1 : ref0) : hit0 ? ref0 : hit1 ? ref1 : TB_REF_POP;
            }
            // 6. triangle step for the parked leaves
            if (do_tri) {
                const int rank = ~pend;
                if (COUNT) ++n_tris;
                float t, u, v;
                if (tb_triangle(r, v3(ta.x, ta.y, ta.z), v3(tb.x, tb.y, tb.z), v3(tc.x, tc.y, tc.z), &t, &u, &v)) {
                    if (t < best_t || (t == best_t && best_rank >= 0 && rank < best_rank)) {
                        best_t = t; best_rank = rank; best_u = u; best_v = v;
                        cull_t = best_t + best_t * 1e-5f;
                    }
                }
                pend = TB_REF_DONE;
            }
            // 7. leave the rounds when the warp is done or has too few rays left to be worth a round
            const unsigned live = __ballot_sync(FULL, cur != TB_REF_DONE || pend != TB_REF_DONE);
            if (live == 0u || (!exhausted && __popc(live) < a.refetch_below)) break;
        }

        // ---- retire finished rays (warp converged)
        const bool finished = cur == TB_REF_DONE && pend == TB_REF_DONE && slot >= 0;
        int preset = -1;
        if (finished) {
            float4 h;
            if (best_rank >= 0) {
                const float4 ta = __ldg(a.tris + (size_t)best_rank * 3);
                const float4 tb = __ldg(a.tris + (size_t)best_rank * 3 + 1);
                h.x = best_t;
                h.y = a.emit_prim ? ta.w : __int_as_float(best_rank);
                h.z = best_u; h.w = best_v;
                preset = (__float_as_int(tb.w) >> 8) & 0xff;
            } else {
                h.x = -1.f; h.y = __int_as_float(-1); h.z = 0.f; h.w = 0.f;
            }
            a.hit_out[slot] = h;
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
        if (finished) slot = -1;
    }
    if (COUNT) {
        atomicAdd(a.visit_counters, (unsigned long long)n_nodes);
        atomicAdd(a.visit_counters + 1, (unsigned long long)n_tris);
    }
}
