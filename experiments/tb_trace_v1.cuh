// tb_trace_v1.cuh — first traversal kernel (if-if: leaves tested inside the node loop). Kept for A/B runs
// (TERRA_B200_TRACE=v1); included by tb_render.cu after the ray/box/triangle helpers.
#pragma once

// Persistent CTAs; each warp pulls rays from a global cursor and refills idle lanes when too few
// are busy (dynamic fetch). Near-first ordered descent of the reference tree with t culling: the
// reference walks every box the ray passes and keeps min t with first-found-wins on ties
// (TerraBVH.c:266-306); culling by the current closest t plus the smallest-rank tie
// rule returns the same primitive. Stack: TB_STACK_SMEM entries per thread in shared memory
// ([depth][thread], conflict free), the rest in local memory.
template <bool COUNT>
__global__ void __launch_bounds__(TB_TRACE_THREADS, TB_TRACE_CTAS_PER_SM) k_trace_v1(TbTraceArgs a) {
    __shared__ int   s_node[TB_STACK_SMEM][TB_TRACE_THREADS];
    __shared__ float s_tmin[TB_STACK_SMEM][TB_TRACE_THREADS];
    int   l_node[TB_STACK_LOCAL];
    float l_tmin[TB_STACK_LOCAL];
    const int tid = threadIdx.x;
    const unsigned lane = tid & 31u;
    const int total = a.count_ptr ? *a.count_ptr : a.count;

    const int DONE = -1;
    int   slot = -1, node = DONE, sp = 0;
    int   best_rank = -1;
    float best_t = FLT_MAX, cull_t = FLT_MAX, best_u = 0.f, best_v = 0.f;
    TbRay r;
    r.o = r.inv = v3s(0.f); r.sx = r.sy = r.sz = 0.f; r.kx = r.ky = r.kz = 0;
    bool exhausted = false;
    unsigned n_nodes = 0, n_tris = 0;        // COUNT variant: the canonical work counters of SURVEY §8(d)

    while (true) {
        // ---- refill idle lanes (warp converged here)
        const bool idle = node == DONE;
        const unsigned idle_mask = __ballot_sync(0xffffffffu, idle);
        if (idle_mask && !exhausted) {
            const int want = __popc(idle_mask);
            int base = 0;
            if (lane == (unsigned)(__ffs(idle_mask) - 1)) base = atomicAdd(a.cursor, want);
            base = __shfl_sync(0xffffffffu, base, __ffs(idle_mask) - 1);
            if (base + want >= total) exhausted = true;
            if (idle) {
                const int q = base + __popc(idle_mask & ((1u << lane) - 1u));
                if (q < total) {
                    slot = a.queue ? a.queue[q] : q;
                    const float4 o4 = a.ray_o[slot], d4 = a.ray_d[slot];
                    r = tb_ray_setup(v3(o4.x, o4.y, o4.z), v3(d4.x, d4.y, d4.z));
                    node = 0; sp = 0;
                    best_rank = -1; best_t = FLT_MAX; cull_t = FLT_MAX;
                }
            }
        }
        if (__ballot_sync(0xffffffffu, node != DONE) == 0u) break;

        // ---- traverse
        while (node != DONE) {
            const float4* np = a.nodes + (size_t)node * 4;
            const float4 q0 = __ldg(np), q1 = __ldg(np + 1), q2 = __ldg(np + 2), q3 = __ldg(np + 3);
            if (COUNT) ++n_nodes;
            const int ref0 = __float_as_int(q3.x), ref1 = __float_as_int(q3.y);
            float tmin0, tmin1;
            bool hit0 = tb_box(r, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, &tmin0);
            bool hit1 = tb_box(r, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, &tmin1);
            hit0 = hit0 && !(tmin0 > cull_t);
            hit1 = hit1 && !(tmin1 > cull_t);
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const int  ref = c ? ref1 : ref0;
                const bool h = c ? hit1 : hit0;
                const float tm = c ? tmin1 : tmin0;
                if (h && ref < 0 && !(tm > cull_t)) {
                    const int rank = ~ref;
                    const float4* tp = a.tris + (size_t)rank * 3;
                    const float4 ta = __ldg(tp), tb = __ldg(tp + 1), tc = __ldg(tp + 2);
                    float t, u, v;
                    if (COUNT) ++n_tris;
                    if (tb_triangle(r, v3(ta.x, ta.y, ta.z), v3(tb.x, tb.y, tb.z), v3(tc.x, tc.y, tc.z), &t, &u, &v)) {
                        if (t < best_t || (t == best_t && best_rank >= 0 && rank < best_rank)) {
                            best_t = t; best_rank = rank; best_u = u; best_v = v;
                            cull_t = best_t + best_t * 1e-5f;
                        }
                    }
                }
            }
            const bool in0 = hit0 && ref0 >= 0 && !(tmin0 > cull_t);
            const bool in1 = hit1 && ref1 >= 0 && !(tmin1 > cull_t);
            if (in0 && in1) {
                const bool near1 = tmin1 < tmin0;
                const int   far_node = near1 ? ref0 : ref1;
                const float far_tmin = near1 ? tmin0 : tmin1;
                node = near1 ? ref1 : ref0;
                if (sp < TB_STACK_SMEM) { s_node[sp][tid] = far_node; s_tmin[sp][tid] = far_tmin; }
                else { l_node[sp - TB_STACK_SMEM] = far_node; l_tmin[sp - TB_STACK_SMEM] = far_tmin; }
                ++sp;
            } else if (in0) {
                node = ref0;
            } else if (in1) {
                node = ref1;
            } else {
                node = DONE;
                while (sp > 0) {
                    --sp;
                    int n; float tm;
                    if (sp < TB_STACK_SMEM) { n = s_node[sp][tid]; tm = s_tmin[sp][tid]; }
                    else { n = l_node[sp - TB_STACK_SMEM]; tm = l_tmin[sp - TB_STACK_SMEM]; }
                    if (!(tm > cull_t)) { node = n; break; }
                }
            }
            if (!exhausted && __popc(__activemask()) < TB_REFETCH_BELOW) break;
        }

        // ---- retire finished rays (warp converged again)
        const bool finished = node == DONE && slot >= 0;
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
                const unsigned m = __ballot_sync(0xffffffffu, finished && preset == p);
                if (m) {
                    int base = 0;
                    if (lane == (unsigned)(__ffs(m) - 1)) base = atomicAdd(a.matq_count + p, __popc(m));
                    base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
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
