// tb_trace_v5.cuh — closest hit for incoherent rays over a 4-WIDE tree: the rounds kernel of tb_trace_v4.cuh
// with a node step that tests four child boxes per fetch.
//
// Why (ncu n3, k_trace_v4 on C3): per ray, 40 rounds for 22 node visits; 44 % of the thread instructions
// are the per-round control code and every round exposes one dependent L1/L2 latency. A 4-wide node
// (tb_bvh.cpp, tb_bvh_collapse4: each node absorbs the grandchildren of its larger children) halves the
// rounds; the box arithmetic per ray stays about the same.
//
// Node, 64 bytes = two 256-bit loads: four child boxes on the global 16-bit grid of the quantised binary
// nodes (same decode: one PRMT per plane, folded into the FMA of the slab test), then four child refs
// (>= 0 node, < 0 leaf ~rank, TB_REF_DONE = empty slot):
//   words 3c .. 3c+2 (c = 0..3): min.x | min.y << 16,  min.z | max.x << 16,  max.y | max.z << 16
//   words 12 .. 15: ref0 .. ref3
// The four (entry distance, ref) pairs are sorted with a 5-comparator network (misses carry +inf); the
// nearest becomes the lane's next node, the others are pushed far-to-near with unconditional stores whose
// stack-pointer increments are the only data-dependent part (no branch).
// Same answers as every other kernel: conservative boxes, the same tb_triangle, the same tie rule.
#pragma once

#define TB_V5_DEPTH 20          /* 3 pushes per node at most; 23 x 128 x 8 B = 23 KB per CTA */

#define TB_V5_CE(TA, RA, TB_, RB)                                                                    \
    do {                                                                                             \
        const bool sw_ = (TB_) < (TA);                                                               \
        const float tl_ = sw_ ? (TB_) : (TA), th_ = sw_ ? (TA) : (TB_);                              \
        const int   rl_ = sw_ ? (RB) : (RA),  rh_ = sw_ ? (RA) : (RB);                               \
        (TA) = tl_; (TB_) = th_; (RA) = rl_; (RB) = rh_;                                             \
    } while (0)

template <bool SPILL>
__global__ void __launch_bounds__(TB_TRACE_THREADS, TB_TRACE_CTAS_PER_SM) k_trace_v5(TbTraceArgs a) {
    __shared__ int2 s_stack[TB_V5_DEPTH + 3][TB_TRACE_THREADS];     // +3: scratch slots of the unconditional pushes
    int2 l_stack[SPILL ? (TB_STACK_SMEM + TB_STACK_LOCAL - TB_V5_DEPTH + 3) : 1];   // host checks the 4-wide tree's stack bound
    const int tid = threadIdx.x;
    const unsigned lane = tid & 31u;
    const unsigned FULL = 0xffffffffu;
    const int total = a.count_ptr ? *a.count_ptr : a.count;
    int2* const my_stack = &s_stack[0][tid];

    int   slot = -1, cur = TB_REF_DONE, pend = TB_REF_DONE, sp = 0;
    int   best_rank = -1;
    float best_t = FLT_MAX, cull_t = FLT_MAX, best_u = 0.f, best_v = 0.f;
    TbRay r;
    r.o = r.inv = v3s(0.f); r.sx = r.sy = r.sz = 0.f; r.kx = r.ky = r.kz = 0;
    TbFastRay fr;
    fr.ix = fr.iy = fr.iz = fr.ox = fr.oy = fr.oz = 0.f;
    bool exhausted = false;
    int  leave_below = a.refetch_below > 1 ? a.refetch_below : 1;

#define TB_V5_PUSH(REF, T)                                                                           \
    do {                                                                                             \
        const int2 e_ = make_int2((REF), __float_as_int(T));                                         \
        if (SPILL && sp >= TB_V5_DEPTH) l_stack[sp - TB_V5_DEPTH] = e_;                              \
        else my_stack[sp * TB_TRACE_THREADS] = e_;                                                   \
        sp += (T) < __int_as_float(0x7f800000) ? 1 : 0;                                              \
    } while (0)

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
                    fr = tb_quant_ray(r, d, a.grid_lo, a.grid_cell);
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
                const bool is_leaf = (unsigned)cur > (unsigned)TB_REF_POP;
                const bool park = is_leaf && pend == TB_REF_DONE;
                pend = park ? cur : pend;
                cur = park ? TB_REF_POP : cur;
            }
            // 2. one pop, resolved with selects (a culled entry costs the lane this round)
            {
                const int top = sp > 0 ? sp - 1 : 0;
                int2 e;
                if (SPILL && top >= TB_V5_DEPTH) e = l_stack[top - TB_V5_DEPTH];
                else e = my_stack[top * TB_TRACE_THREADS];
                const bool need = cur == TB_REF_POP;
                const bool empty = sp == 0;
                const int popped = __int_as_float(e.y) > cull_t ? TB_REF_POP : e.x;
                cur = need ? (empty ? TB_REF_DONE : popped) : cur;
                sp = (need && !empty) ? top : sp;
            }
            // 3. what does this round do?
            const bool is_node = cur >= 0;
            const unsigned node_mask = __ballot_sync(FULL, is_node);
            const unsigned pend_mask = __ballot_sync(FULL, pend != TB_REF_DONE);
            const bool tri_round = pend_mask != 0u && (__popc(pend_mask) >= a.tri_min || node_mask == 0u);
            const bool do_tri = tri_round && pend != TB_REF_DONE;
            // 4. all loads of the round first
            uint4 qa, qb, qc, qd;
            float4 ta, tb, tc;
            if (is_node) {
                const uint4* np = a.qnodes4 + (size_t)cur * 4;
                asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                             : "=r"(qa.x), "=r"(qa.y), "=r"(qa.z), "=r"(qa.w), "=r"(qb.x), "=r"(qb.y), "=r"(qb.z), "=r"(qb.w) : "l"(np));
                asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                             : "=r"(qc.x), "=r"(qc.y), "=r"(qc.z), "=r"(qc.w), "=r"(qd.x), "=r"(qd.y), "=r"(qd.z), "=r"(qd.w) : "l"(np + 2));
            }
            if (do_tri) {
                const float4* tp = a.tris + (size_t)(~pend) * 3;
                ta = __ldg(tp); tb = __ldg(tp + 1); tc = __ldg(tp + 2);
            }
            // 5. node step: four boxes, sort, descend into the nearest, push the rest far-to-near
            if (is_node) {
                const float INF = __int_as_float(0x7f800000);
                float t0, t1, t2, t3;
                int r0 = (int)qd.x, r1 = (int)qd.y, r2 = (int)qd.z, r3 = (int)qd.w;
                const bool h0 = tb_box_fast(fr, tb_qlo(qa.x), tb_qhi(qa.x), tb_qlo(qa.y), tb_qhi(qa.y), tb_qlo(qa.z), tb_qhi(qa.z), &t0);
                const bool h1 = tb_box_fast(fr, tb_qlo(qa.w), tb_qhi(qa.w), tb_qlo(qb.x), tb_qhi(qb.x), tb_qlo(qb.y), tb_qhi(qb.y), &t1);
                const bool h2 = tb_box_fast(fr, tb_qlo(qb.z), tb_qhi(qb.z), tb_qlo(qb.w), tb_qhi(qb.w), tb_qlo(qc.x), tb_qhi(qc.x), &t2);
                const bool h3 = tb_box_fast(fr, tb_qlo(qc.y), tb_qhi(qc.y), tb_qlo(qc.z), tb_qhi(qc.z), tb_qlo(qc.w), tb_qhi(qc.w), &t3);
                t0 = (h0 && !(t0 > cull_t)) ? t0 : INF;                    // slots 0 and 1 are never empty
                t1 = (h1 && !(t1 > cull_t)) ? t1 : INF;
                t2 = (h2 && !(t2 > cull_t) && r2 != TB_REF_DONE) ? t2 : INF;
                t3 = (h3 && !(t3 > cull_t) && r3 != TB_REF_DONE) ? t3 : INF;
                TB_V5_CE(t0, r0, t1, r1);
                TB_V5_CE(t2, r2, t3, r3);
                TB_V5_CE(t0, r0, t2, r2);
                TB_V5_CE(t1, r1, t3, r3);
                TB_V5_CE(t1, r1, t2, r2);
                TB_V5_PUSH(r3, t3);
                TB_V5_PUSH(r2, t2);
                TB_V5_PUSH(r1, t1);
                cur = t0 < INF ? r0 : TB_REF_POP;
            }
            // 6. triangle step for the parked leaves
            if (do_tri) {
                const int rank = ~pend;
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
#undef TB_V5_PUSH
}
