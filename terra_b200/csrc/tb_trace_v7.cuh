// tb_trace_v7.cuh — closest hit for INCOHERENT rays on the 4-WIDE tree (tb_bvh4_build): k_trace_v6's rounds with four
// child boxes per node step.
//
// Why. k_trace_v6 is bound by issue slots and node-fetch latency together (profiles/r2_summary.md): a round issues
// ~87 warp instructions behind one dependent 32-byte fetch, and a ray needs 22 such fetches on the 1M-triangle
// terrain. A 4-wide node is one 64-byte fetch (two 256-bit loads of one line pair) behind which a round does twice
// the box work, and a ray needs about half as many of them: the same ALU work per ray in half as many latency-exposed
// steps, with half the per-round bookkeeping.
//
// Same answer as every other trace kernel: the child boxes are the binary tree's own quantised boxes (copied word for
// word), the leaves, ranks and triangle records are the binary tree's, the triangle test is tb_triangle, equal t goes
// to the smaller rank. Only the order in which boxes are met changes.
//
// Stack: a node step pushes up to three entries, so the bound is the sum of (children - 1) along a path
// (TbQuadBvh::stack_bound, 30-40 for a million triangles): the shared rows hold the usual occupancy and the SPILL
// instantiation's local array the rest.
#pragma once

__device__ __forceinline__ bool tb_box_sel_cull(const TbSelRay& f, unsigned wx, unsigned wy, unsigned wz, float cull_t, float* tmin_out) {
    const float ax = __fmaf_rn(tb_qsel(wx, f.nx), f.ix, f.ox), bx = __fmaf_rn(tb_qsel(wx, f.fx), f.ix, f.ox);
    const float ay = __fmaf_rn(tb_qsel(wy, f.ny), f.iy, f.oy), by = __fmaf_rn(tb_qsel(wy, f.fy), f.iy, f.oy);
    const float az = __fmaf_rn(tb_qsel(wz, f.nz), f.iz, f.oz), bz = __fmaf_rn(tb_qsel(wz, f.fz), f.iz, f.oz);
    const float tmin = fmaxf(fmaxf(fmaxf(ax, ay), az), 0.f);
    const float tmax = fminf(fminf(fminf(bx, by), bz), cull_t);
    *tmin_out = tmin;
    return tmax > tmin;
}

__device__ __forceinline__ void tb_sts64_if(bool p, unsigned addr, int2 v) {
    asm volatile("{ .reg .pred q; setp.ne.u32 q, %3, 0; @q st.shared.v2.b32 [%0], {%1, %2}; }" :: "r"(addr), "r"(v.x), "r"(v.y), "r"((unsigned)p) : "memory");
}

// Pop for the 4-wide kernel: the entry carries its box's entry distance; one that the closest hit has culled since it
// was pushed leaves the lane in POP state (it pops again at the next step).
template <bool SPILL, int ROW>
__device__ __forceinline__ void tb_v7_pop(int& cur, unsigned& sp, float cull_t, unsigned smem_end, const int2* l_stack) {
    int2 e;
    if (SPILL && sp > smem_end) e = l_stack[(sp - smem_end) / ROW - 1];
    else e = tb_lds64(sp - ROW);
    const bool need = cur == TB_REF_POP;
    const int popped = __int_as_float(e.y) > cull_t ? TB_REF_POP : e.x;
    cur = need ? popped : cur;
    sp = need ? sp - ROW : sp;
}

// Node step: four boxes, the hit children ordered by entry distance with a 5-comparator network on (key, ref) pairs —
// key = the bits of tmin (>= 0: they order as unsigned) with the slot number in the two low bits, so keys are distinct
// and a stored tmin is rounded DOWN (culls conservatively) — the nearest becomes the lane's node, the others are pushed
// far to near.
template <bool SPILL, int ROW>
__device__ __forceinline__ void tb_v7_node_step(const TbSelRay& sr, const uint4& qa, const uint4& qb, const uint4& qc, const uint4& qd, float cull_t,
                                                int& cur, unsigned& sp, unsigned smem_end, int2* l_stack) {
    float t0, t1, t2, t3;
    // the cull distance is folded into the exit distance (a box entered at or beyond cull_t holds no closer hit)
    const bool h0 = tb_box_sel_cull(sr, qa.x, qa.y, qa.z, cull_t, &t0);
    const bool h1 = tb_box_sel_cull(sr, qa.w, qb.x, qb.y, cull_t, &t1);
    const bool h2 = tb_box_sel_cull(sr, qb.z, qb.w, qc.x, cull_t, &t2);
    const bool h3 = tb_box_sel_cull(sr, qc.y, qc.z, qc.w, cull_t, &t3);
    unsigned k0 = h0 ? (__float_as_uint(t0) & ~3u) : 0xffffffffu;
    unsigned k1 = h1 ? ((__float_as_uint(t1) & ~3u) | 1u) : 0xffffffffu;
    unsigned k2 = h2 ? ((__float_as_uint(t2) & ~3u) | 2u) : 0xffffffffu;
    unsigned k3 = h3 ? ((__float_as_uint(t3) & ~3u) | 3u) : 0xffffffffu;
    int r0 = (int)qd.x, r1 = (int)qd.y, r2 = (int)qd.z, r3 = (int)qd.w;
#define TB_CSWAP(KA, RA, KB, RB) { const bool s_ = KA > KB; const unsigned kl_ = s_ ? KB : KA, kh_ = s_ ? KA : KB; \
                       const int rl_ = s_ ? RB : RA, rh_ = s_ ? RA : RB; KA = kl_; KB = kh_; RA = rl_; RB = rh_; }
    TB_CSWAP(k0, r0, k1, r1) TB_CSWAP(k2, r2, k3, r3) TB_CSWAP(k0, r0, k2, r2) TB_CSWAP(k1, r1, k3, r3) TB_CSWAP(k1, r1, k2, r2)
#undef TB_CSWAP
    // misses sort to the end: entry j is a hit iff kj != ~0. The key itself is stored as the entry's distance:
    // at most 3 ulp above the box's tmin, far inside the 1e-5 margin of cull_t.
    if (!SPILL || sp + 2 * ROW < smem_end) {
        tb_sts64_if(k3 != 0xffffffffu, sp, make_int2(r3, (int)k3)); sp += k3 != 0xffffffffu ? ROW : 0;
        tb_sts64_if(k2 != 0xffffffffu, sp, make_int2(r2, (int)k2)); sp += k2 != 0xffffffffu ? ROW : 0;
        tb_sts64_if(k1 != 0xffffffffu, sp, make_int2(r1, (int)k1)); sp += k1 != 0xffffffffu ? ROW : 0;
    } else {                        // near the end of the shared rows: entries from smem_end on live in local memory
        if (k3 != 0xffffffffu) {
            if (sp >= smem_end) l_stack[(sp - smem_end) / ROW] = make_int2(r3, (int)k3); else tb_sts64(sp, make_int2(r3, (int)k3));
            sp += ROW;
        }
        if (k2 != 0xffffffffu) {
            if (sp >= smem_end) l_stack[(sp - smem_end) / ROW] = make_int2(r2, (int)k2); else tb_sts64(sp, make_int2(r2, (int)k2));
            sp += ROW;
        }
        if (k1 != 0xffffffffu) {
            if (sp >= smem_end) l_stack[(sp - smem_end) / ROW] = make_int2(r1, (int)k1); else tb_sts64(sp, make_int2(r1, (int)k1));
            sp += ROW;
        }
    }
    cur = k0 != 0xffffffffu ? r0 : TB_REF_POP;
}

#ifndef TB_V7_STEPS
#define TB_V7_STEPS 2           /* node steps per round (see tb_trace_v8.cuh, step 6b): 1 / 2 / 3 = 4127 / 4243 / 4180 Mrays/s on C3 10M (profiles/r3b) */
#endif
#ifndef TB_V7_CTAS_PER_SM
#define TB_V7_CTAS_PER_SM 7     /* 72 registers, no spills (the ray state a round does not read lives in shared memory, as in
                                   tb_trace_v8.cuh); +6.3 % over 6 CTAs x 80 registers on the 10M-triangle terrain (profiles/r2v) */
#endif

template <bool COUNT, bool SPILL>
__global__ void __launch_bounds__(TB_TRACE_THREADS, TB_V7_CTAS_PER_SM) k_trace_v7(TbTraceArgs a) {
    extern __shared__ __align__(16) unsigned char s_v6_stack[];    // [stack_smem + 3][threads] x 8 B entries, [TB_LEAN_WORDS][threads] words, [warps] x 8 B refill state
    constexpr int ROW = TB_TRACE_THREADS * 8;
    int2 l_stack[SPILL ? TB_V7_LOCAL : 1];
    const int tid = threadIdx.x;
    const unsigned lane = tid & 31u;
    const unsigned FULL = 0xffffffffu;
    const int total = a.count_ptr ? *a.count_ptr : a.count;
    // entry k of this lane's stack lives at bottom + k * ROW (shared-space addresses); the row below entry 0 holds the
    // sentinel an empty pop reads, the row above the deepest entry is the scratch an unconditional push writes
    // (one more row below the sentinel: the pop is unconditional, and after the sentinel has been popped the lanes
    // that are finished keep reading one row further down until the warp refills)
    const unsigned bottom = (unsigned)__cvta_generic_to_shared(s_v6_stack) + 2 * ROW + tid * 8;
    tb_sts64(bottom - ROW, make_int2(TB_REF_DONE, 0));             // never culled (tmin 0), never overwritten
    const unsigned smem_end = bottom + a.stack_smem * ROW;         // SPILL: entries from this address on live in local memory

    // what only a triangle round or the retirement of a ray reads lives in shared memory, [word][thread] behind the
    // stack rows (see tb_trace_v8.cuh): 0-2 offset origin, 3-5 shear constants, 6 kx | ky << 2 | kz << 4, 7 best rank,
    // 8 best tag, 9 slot
    constexpr int LW = TB_TRACE_THREADS * 4;
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
            // 1. park a leaf (a further leaf waits in `cur` until a parking slot is free)
            {
                const bool is_leaf = (unsigned)cur > (unsigned)TB_REF_POP;       // ~rank: 0x80000002 .. 0xffffffff
                const bool park = is_leaf && pend == TB_REF_DONE;
                pend = park ? cur : pend;
                cur = park ? TB_REF_POP : cur;
            }
            // 2. pop, resolved with selects (a second pop in the same round, and a pop loop, were measured: -1.5 % and -10 %,
            //    experiments/README.md)
            tb_v7_pop<SPILL, ROW>(cur, sp, cull_t, smem_end, l_stack);
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
                    if (cur == TB_REF_POP) ++n_culled;                                   // popped an entry the closest hit had culled
                    else if (cur != TB_REF_DONE) ++n_second;                             // a leaf, while the parked one waits
                    else if (pend != TB_REF_DONE) ++n_parked;
                    else ++n_empty;
                }
            }
            // 4. all loads of the round first, so node and triangle latencies overlap
            float4 ta, tb, tc;
            uint4 qa, qb, qc, qd;                       // 16 words: boxes of children 0..3 (3 words each), then the four refs
            if (is_node) {
                tb_ld_qnode(a.q4nodes + (size_t)cur * 4, qa, qb);
                tb_ld_qnode(a.q4nodes + (size_t)cur * 4 + 2, qc, qd);
            }
            if (do_tri) {
                const float4* tp = a.tris + (size_t)(~pend) * 3;
                ta = __ldg(tp); tb = __ldg(tp + 1); tc = __ldg(tp + 2);
            }
            // 5. node step
            if (is_node) {
                if (COUNT) ++n_nodes;
                tb_v7_node_step<SPILL, ROW>(sr, qa, qb, qc, qd, cull_t, cur, sp, smem_end, l_stack);
            }
            // 6. triangle step for the parked leaves
            if (do_tri) {
                const int rank = ~pend;
                if (COUNT) ++n_tris;
                float t, u, v;
                TbRay r;
                r.o = v3(__int_as_float(tb_lds32(lean + 0 * LW)), __int_as_float(tb_lds32(lean + 1 * LW)), __int_as_float(tb_lds32(lean + 2 * LW)));
                r.inv = v3s(0.f);
                r.sx = __int_as_float(tb_lds32(lean + 3 * LW)); r.sy = __int_as_float(tb_lds32(lean + 4 * LW)); r.sz = __int_as_float(tb_lds32(lean + 5 * LW));
                const int kpack = tb_lds32(lean + 6 * LW);
                r.kx = kpack & 3; r.ky = (kpack >> 2) & 3; r.kz = (kpack >> 4) & 3;
                if (tb_triangle(r, v3(ta.x, ta.y, ta.z), v3(tb.x, tb.y, tb.z), v3(tc.x, tc.y, tc.z), &t, &u, &v)) {
                    bool better = t < best_t;
                    if (t == best_t) {                                           // exact tie: the smaller rank wins
                        const int best_rank = tb_lds32(lean + 7 * LW);
                        better = best_rank >= 0 && rank < best_rank;
                    }
                    if (better) {
                        best_t = t;
                        tb_sts32(lean + 7 * LW, rank); tb_sts32(lean + 8 * LW, __float_as_int(tb.w));
                        // API rays: object | triangle << 8, u, v; wavefront: rank and the object | preset << 8 tag for the
                        // shade pass (u, v are not used there). A later, closer hit overwrites the record.
                        const int slot = tb_lds32(lean + 9 * LW);
                        tb_st_stream(a.hit_out + slot, make_float4(t, a.emit_prim ? ta.w : __int_as_float(rank), a.emit_prim ? u : tb.w, v));
                    }
                }
                pend = TB_REF_DONE;
            }
            // 6b. further node steps of the round (tb_trace_v8.cuh): the round's bookkeeping is paid once for TB_V7_STEPS steps
#pragma unroll
            for (int step = 1; step < TB_V7_STEPS; ++step) {
                const float cull2 = best_t + best_t * 1e-5f;                 // the triangle step may have moved the closest hit
                tb_v7_pop<SPILL, ROW>(cur, sp, cull2, smem_end, l_stack);
                const bool is_node2 = cur >= 0;
                if (COUNT) {
                    ++n_rounds;
                    if (!is_node2) {
                        if (cur == TB_REF_POP) ++n_culled;
                        else if (cur != TB_REF_DONE) ++n_second;
                        else if (pend != TB_REF_DONE) ++n_parked;
                        else ++n_empty;
                    }
                }
                if (is_node2) {
                    uint4 qe, qf, qg, qh;
                    tb_ld_qnode(a.q4nodes + (size_t)cur * 4, qe, qf);
                    tb_ld_qnode(a.q4nodes + (size_t)cur * 4 + 2, qg, qh);
                    if (COUNT) ++n_nodes;
                    tb_v7_node_step<SPILL, ROW>(sr, qe, qf, qg, qh, cull2, cur, sp, smem_end, l_stack);
                }
            }
            // 7. leave the rounds when the warp is done or has too few rays left to be worth a round
            const unsigned live = __ballot_sync(FULL, cur != TB_REF_DONE || pend != TB_REF_DONE);
            if (live == 0u || (!drained && __popc(live) < a.refetch_below)) break;
        }

        // ---- retire finished rays (warp converged)
        // a ray ends by popping the sentinel below entry 0 (sp == bottom - ROW); a lane without a ray has sp == bottom
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
