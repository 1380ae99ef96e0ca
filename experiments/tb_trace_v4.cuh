// tb_trace_v4.cuh — closest-hit kernel, rounds structure (tb_trace_v3.cuh) with the control path rebuilt.
//
// Why (ncu r1b, source page of k_trace_v3 on C2 bounce rays): 59 % of all warp instructions were the
// round's CONTROL code, not boxes (22 %) or triangles (16 %), and the worst lines ran nearly empty —
// the push under `if (hit0 && hit1)` with 3.9 of 32 lanes, the pop under `if (cur == POP)` with 3.9,
// the smem/local split of both with its own branches. A divergent block costs its full instruction
// count however few lanes take it. Here:
//   - a stack entry is ONE 64-bit word {ref, tmin} in shared memory, [depth][thread] (conflict-free):
//     one STS.64 / LDS.64 instead of two 32-bit accesses to two arrays;
//   - the push is unconditional for every lane that tested a node (the slot above the top is scratch)
//     and only `sp += both_hit` is data dependent: no branch;
//   - the pop reads the top entry unconditionally and resolves the state with selects: no branch;
//   - a leaf is parked BEFORE the pop, so the lane pops and fetches its next node in the same round;
//   - the local-memory spill exists only in the SPILL instantiation, chosen by the host when the tree
//     is deeper than the shared stack (TB_V4_DEPTH entries; C2 needs 21, the 1M-triangle terrain 24: the
//     spill branch is there but almost never taken, near-first order keeps the stack shallow).
// Results are identical to k_trace_v3 (same boxes, same triangle test, same tie rule).
#pragma once

#define TB_V4_DEPTH 16          /* 17 x 128 x 8 B = 17 KB per CTA: the rest of the SM's 256 KB stays L1 for the nodes */

// OPT bit 0: fetch the 32-byte quantised node with ONE 256-bit load (LDG.E.256, sm_100) instead of two
//            128-bit ones — a divergent warp pays one L1 tag wavefront per lane per load instruction;
// OPT bit 1: single pop per round (a culled entry costs the lane the round) instead of the pop loop.
// (10 CTAs per SM at 51 registers was tried: 112 bytes of spills, -2 % on C3.)
template <bool COUNT, bool EXACT, bool QUANT, bool SPILL, int OPT>
__global__ void __launch_bounds__(TB_TRACE_THREADS, TB_TRACE_CTAS_PER_SM) k_trace_v4(TbTraceArgs a) {
    __shared__ int2 s_stack[TB_V4_DEPTH + 1][TB_TRACE_THREADS];     // +1: scratch slot for the unconditional push
    int2 l_stack[SPILL ? (TB_STACK_SMEM + TB_STACK_LOCAL - TB_V4_DEPTH + 1) : 1];
    const int tid = threadIdx.x;
    const unsigned lane = tid & 31u;
    const unsigned FULL = 0xffffffffu;
    const int total = a.count_ptr ? *a.count_ptr : a.count;
    int2* const my_stack = &s_stack[0][tid];                        // entry k at my_stack[k * TB_TRACE_THREADS]

    int   slot = -1, cur = TB_REF_DONE, pend = TB_REF_DONE, sp = 0;
    int   best_rank = -1;
    float best_t = FLT_MAX, cull_t = FLT_MAX, best_u = 0.f, best_v = 0.f;
    TbRay r;
    r.o = r.inv = v3s(0.f); r.sx = r.sy = r.sz = 0.f; r.kx = r.ky = r.kz = 0;
    TbFastRay fr;
    fr.ix = fr.iy = fr.iz = fr.ox = fr.oy = fr.oz = 0.f;
    bool exhausted = false;
    int  leave_below = a.refetch_below > 1 ? a.refetch_below : 1;                           // 1 once the ray queue is empty
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
            if (base + want >= total) { exhausted = true; leave_below = 1; }
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
            // 1. park a leaf (a second leaf waits in `cur` until the parked one has been tested)
            {
                const bool is_leaf = (unsigned)cur > (unsigned)TB_REF_POP;       // ~rank: 0x80000002 .. 0xffffffff
                const bool park = is_leaf && pend == TB_REF_DONE;
                pend = park ? cur : pend;
                cur = park ? TB_REF_POP : cur;
            }
            // 2. pop, resolved with selects; entries the closest hit has culled since they were pushed are
            //    skipped in a short warp-wide loop (7 instructions a turn) instead of costing the lane a round
            do {
                const int top = sp > 0 ? sp - 1 : 0;
                int2 e;
                if (SPILL && top >= TB_V4_DEPTH) e = l_stack[top - TB_V4_DEPTH];
                else e = my_stack[top * TB_TRACE_THREADS];
                const bool need = cur == TB_REF_POP;
                const bool empty = sp == 0;
                const int popped = __int_as_float(e.y) > cull_t ? TB_REF_POP : e.x;
                cur = need ? (empty ? TB_REF_DONE : popped) : cur;
                sp = (need && !empty) ? top : sp;
            } while (!(OPT & 2) && __any_sync(FULL, cur == TB_REF_POP));
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
                    if (OPT & 1) {
                        asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                     : "=r"(qa.x), "=r"(qa.y), "=r"(qa.z), "=r"(qa.w), "=r"(qb.x), "=r"(qb.y), "=r"(qb.z), "=r"(qb.w) : "l"(np));
                    } else { qa = __ldg(np); qb = __ldg(np + 1); }
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
                    hit0 = tb_box_fast(fr, tb_qlo(qa.x), tb_qlo(qa.y), tb_qlo(qa.z), tb_qhi(qa.x), tb_qhi(qa.y), tb_qhi(qa.z), &tmin0);
                    hit1 = tb_box_fast(fr, tb_qlo(qa.w), tb_qlo(qb.x), tb_qlo(qb.y), tb_qhi(qa.w), tb_qhi(qb.x), tb_qhi(qb.y), &tmin1);
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
                const bool both = hit0 && hit1;
                const int2 far = make_int2(near1 ? ref0 : ref1, __float_as_int(near1 ? tmin0 : tmin1));
                if (SPILL && sp >= TB_V4_DEPTH) l_stack[sp - TB_V4_DEPTH] = far;
                else my_stack[sp * TB_TRACE_THREADS] = far;                 // scratch when !both
                sp += both ? 1 : 0;
                const int one = hit0 ? ref0 : (hit1 ? ref1 : TB_REF_POP);
                cur = both ? (near1 ? ref1 : ref0) : one;
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
            if (__popc(live) < leave_below) break;
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
                h.z = a.emit_prim ? best_u : tb.w;         // wavefront: object | preset << 8 for the shade pass (u, v are not used there)
                    h.w = best_v;
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
