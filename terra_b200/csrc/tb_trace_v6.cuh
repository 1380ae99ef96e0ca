// tb_trace_v6.cuh — closest hit for INCOHERENT rays (bounce, visibility and API rays): the rounds kernel.
//
// What it computes: terra_scene_raycast without the surface record (Terra.c:1623-1657 -> terra_bvh_traverse
// TerraBVH.c:250-310 -> box Terra.c:851-878 -> triangle TerraGeometry.c:159-259), same answer as every other
// trace kernel of this library: same triangles, same tb_triangle, equal t -> smaller rank.
//
// Shape (kept from round 1's k_trace_v4): persistent CTAs, per-lane rays in warp-synchronous ROUNDS — every lane
// that owns a ray fetches one 32-byte quantised node per round with ONE 256-bit load, a leaf is parked and the
// watertight triangle test runs for all parked leaves at once when enough lanes hold one, idle lanes are refilled
// by ballot, the stack is one 64-bit {ref, tmin} word per entry in shared memory, [depth][thread].
//
// What round 2 changed, and why. ncu r1r (k_trace_v4 on C2 bounce rays) showed sm__pipe_alu_cycles_active 70 % against
// sm__pipe_fma 19 % and issue-active 69.5 %: the busiest unit was the half-rate ALU pipe (PRMT / FMNMX / ISETP / SEL),
// 69 of the 146 instructions of a round. Cutting those (first three items) took 20 % of the instructions out
// (185 -> 148 warp instructions per ray on C3 1M, ncu r2d) but only 2 % of the time at equal occupancy and stack
// settings: with the instruction count down, the kernel waits on the node fetch instead (35 % of all stall samples
// sit on the first use of the node load, issue-active 60 %). The rest of the gain came from registers, L1 and L2
// (last items). Measured end to end: 4.29 -> 4.94 Grays/s on C3 1M, profiles/r2_summary.md.
//   - SIGN-SELECTED PLANES. The two planes of an axis share a word of the node (tb_trace.cuh), and the ray carries
//     one byte-permute selector per axis and side, fixed by the sign of its direction: the PRMT that decodes the
//     16-bit grid value picks the near or the far plane directly. The six min / max pairs per box of the generic
//     slab test disappear: a box is 6 PRMT + 6 FFMA + max3, max(., 0), min3 + 2 compares (11 ALU-pipe instructions,
//     was 17).
//   - the stack pointer is the shared-memory ADDRESS of the lane's next free entry, and the row below entry 0 holds
//     a {DONE, 0} sentinel: a pop is one LDS at [sp - row] and three selects, an empty stack pops "finished" by
//     itself, a push is one STS at [sp] — no index arithmetic, no empty test;
//   - the shared stack is sized from the tree (dynamic shared memory, TbTraceArgs::stack_smem rows): trees up to 24
//     levels (C1, C2, C5, C3 at 1M) run without the SPILL code; deeper ones keep 16 rows + local memory, because the
//     shared stack comes out of the same 256 KB as L1 and a 30-row stack at 7 CTAs leaves the node fetches 18 KB of it
//     (r2f A/B on the 10M-triangle terrain: 2.14 Grays/s with 30 rows, 3.23 with 16 + spill);
//   - L2 policy per stream (ncu r2d, C3 1M: L2 hit rate 74 %, 115 B of DRAM traffic per ray — the 32 B rays and 16 B hit
//     records of a 64M-ray batch stream through L2 and push the 80 MB of nodes + triangles out): node loads ask L2 to
//     evict them last, ray loads and hit stores go through an evict-first policy (createpolicy: the descriptor lives
//     in uniform registers, no per-thread cost). +2 % on C3 1M;
//   - the three reciprocals behind the box planes are rcp.approx (1 ulp: four orders of magnitude inside the box
//     padding); the one behind the triangle test stays IEEE (bit-exact t, u, v);
//   - a hit record is written when the closest hit improves (the triangle's tag words are in registers right then),
//     not at retirement: no dependent triangle re-fetch when a ray ends, and u, v need no registers;
//   - registers: 7 CTAs per SM at 72 registers instead of 8 at 64. At 64 the compiler spilled a selector and the
//     refill flag into local memory and re-read them every round (+9 % on C3 1M for the 7-CTA build, r2a A/B); the
//     axis permutation of the triangle test is one packed word and the cull distance is recomputed from the closest
//     hit instead of carried, which is what keeps it at 72;
//   - finished rays are retired only when the warp has one.
#pragma once

#define TB_V6_MAX_SMEM_DEPTH   24          /* (24 + 3) rows x 1 KB x 7 CTAs = 189 KB of the SM's 256 KB of L1 + shared memory */
#define TB_V6_SPILL_SMEM_DEPTH 16
#define TB_V6_LOCAL            (TB_STACK_SMEM + TB_STACK_LOCAL)

struct TbSelRay {              // plane = fma(8388608 + q, i, o), see tb_quant_ray
    float    ix, iy, iz, ox, oy, oz;
    unsigned nx, ny, nz;       // PRMT selectors of the near plane of each axis (0x7610: low half, 0x7632: high half)
    unsigned fx, fy, fz;       // ... of the far plane
};

__device__ __forceinline__ float tb_rcp_approx(float a) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
    return r;
}
__device__ __forceinline__ float tb_finite_inverse_approx(float d) {
    const float lim = 1e-30f;
    const float a = fabsf(d) < lim ? copysignf(lim, d) : d;
    return tb_rcp_approx(a);
}

__device__ __forceinline__ TbSelRay tb_sel_ray(const TbRay& r, V3 d, const float* glo, const float* gcell) {
    TbSelRay f;
    const float ix = tb_finite_inverse_approx(d.x), iy = tb_finite_inverse_approx(d.y), iz = tb_finite_inverse_approx(d.z);
    f.ix = gcell[0] * ix; f.iy = gcell[1] * iy; f.iz = gcell[2] * iz;
    f.ox = __fmaf_rn(-8388608.f, f.ix, (glo[0] - r.o.x) * ix);
    f.oy = __fmaf_rn(-8388608.f, f.iy, (glo[1] - r.o.y) * iy);
    f.oz = __fmaf_rn(-8388608.f, f.iz, (glo[2] - r.o.z) * iz);
    // direction >= 0: planes are entered at min (low half) and left at max (high half); < 0: the other way round
    f.nx = ix < 0.f ? 0x7632u : 0x7610u; f.fx = ix < 0.f ? 0x7610u : 0x7632u;
    f.ny = iy < 0.f ? 0x7632u : 0x7610u; f.fy = iy < 0.f ? 0x7610u : 0x7632u;
    f.nz = iz < 0.f ? 0x7632u : 0x7610u; f.fz = iz < 0.f ? 0x7610u : 0x7632u;
    return f;
}

// prmt.b32 straight (no __byte_perm: with a run-time selector that intrinsic masks it with 0x7777 first, one more
// ALU-pipe instruction per plane)
__device__ __forceinline__ float tb_qsel(unsigned w, unsigned sel) {
    float r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=f"(r) : "r"(w), "r"(0x4B000000u), "r"(sel));
    return r;
}

// One child box, planes already ordered by the ray's direction signs. tmin is clamped at 0 like every other
// box test here (tmax > max(tmin, 0), Terra.c:877).
__device__ __forceinline__ bool tb_box_sel(const TbSelRay& f, unsigned wx, unsigned wy, unsigned wz, float* tmin_out) {
    const float ax = __fmaf_rn(tb_qsel(wx, f.nx), f.ix, f.ox), bx = __fmaf_rn(tb_qsel(wx, f.fx), f.ix, f.ox);
    const float ay = __fmaf_rn(tb_qsel(wy, f.ny), f.iy, f.oy), by = __fmaf_rn(tb_qsel(wy, f.fy), f.iy, f.oy);
    const float az = __fmaf_rn(tb_qsel(wz, f.nz), f.iz, f.oz), bz = __fmaf_rn(tb_qsel(wz, f.fz), f.iz, f.oz);
    const float tmin = fmaxf(fmaxf(fmaxf(ax, ay), az), 0.f);
    const float tmax = fminf(fminf(bx, by), bz);
    *tmin_out = tmin;
    return tmax > tmin;
}

__device__ __forceinline__ int2 tb_lds64(unsigned addr) {
    int2 v;
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void tb_sts64(unsigned addr, int2 v) {
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" :: "r"(addr), "r"(v.x), "r"(v.y) : "memory");
}

__device__ __forceinline__ int tb_lds32(unsigned addr) {
    int v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void tb_sts32(unsigned addr, int v) {
    asm volatile("st.shared.b32 [%0], %1;" :: "r"(addr), "r"(v) : "memory");
}

// Refill of the lean rounds kernels (k_trace_v7, k_trace_v8). Rays are handed out in blocks of TB_REFILL_BLOCK queue
// positions per warp, and the block after the one in use is reserved one refill ahead: the atomic's round trip (2.4 % of
// all stall samples sat on the shuffle behind it, ncu r2y) is over long before its result is read, and a refill that
// still finds enough positions in its block issues no atomic at all (+1.6 % on C3 1M, profiles/r3d; an L2 prefetch of
// the block's rays on top of it measured nothing). {next, end} of the block in use live in 8 bytes of shared memory per
// warp (`wstate`), the position reserved ahead in lane 0's `pre`. Returns the calling lane's queue position (lanes of
// idle_mask only; >= total: no ray left for it) and sets `drained` when this warp will find no further ray.
#define TB_REFILL_BLOCK 32      /* >= 32: a refill takes at most one block switch */
__device__ __forceinline__ int tb_refill_init(unsigned lane, unsigned wstate, int* cursor) {
    int pre = 0;
    if (lane == 0) {
        tb_sts64(wstate, make_int2(0, 0));
        pre = atomicAdd(cursor, TB_REFILL_BLOCK);
    }
    __syncwarp();
    return pre;
}
__device__ __forceinline__ int tb_refill_position(unsigned idle_mask, unsigned lane, unsigned wstate, int& pre, int* cursor, int total, bool& drained) {
    const int want = __popc(idle_mask);
    const int2 blk = tb_lds64(wstate);
    const int avail = blk.y - blk.x;
    const int rank = __popc(idle_mask & ((1u << lane) - 1u));
    int q = blk.x + rank;
    if (want > avail) {                                             // warp-uniform: the block in use runs out
        const int nb = __shfl_sync(0xffffffffu, pre, 0);
        q = rank < avail ? q : nb + (rank - avail);
        if (lane == 0) {
            tb_sts64(wstate, make_int2(nb + (want - avail), nb + TB_REFILL_BLOCK));
            pre = atomicAdd(cursor, TB_REFILL_BLOCK);
        }
        if (nb + (want - avail) >= total) drained = true;          // blocks are handed out in increasing order
    } else {
        if (lane == 0) tb_sts32(wstate, blk.x + want);
        if (blk.x + want >= total) drained = true;
    }
    __syncwarp();
    return q;
}

__device__ __forceinline__ void tb_ld_qnode(const uint4* np, uint4& qa, uint4& qb) {
    asm volatile("ld.global.nc.L2::evict_last.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(qa.x), "=r"(qa.y), "=r"(qa.z), "=r"(qa.w), "=r"(qb.x), "=r"(qb.y), "=r"(qb.z), "=r"(qb.w) : "l"(np));
}
// ray records and hit records: streamed once through the whole hierarchy
__device__ __forceinline__ float4 tb_ld_stream(const float4* p) {
    float4 v;
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ void tb_st_stream(float4* p, float4 v) {
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "l"(pol) : "memory");
}

#ifndef TB_V6_CTAS_PER_SM
#define TB_V6_CTAS_PER_SM 7     /* 72 registers: no spills; measured +9 % over 8 CTAs x 64 registers on C3 1M (profiles/r2a) */
#endif
template <bool COUNT, bool EXACT, bool SPILL>
__global__ void __launch_bounds__(TB_TRACE_THREADS, TB_V6_CTAS_PER_SM) k_trace_v6(TbTraceArgs a) {
    extern __shared__ __align__(16) unsigned char s_v6_stack[];    // [stack_smem + 3][TB_TRACE_THREADS] x 8 B
    constexpr int ROW = TB_TRACE_THREADS * 8;
    int2 l_stack[SPILL ? TB_V6_LOCAL : 1];
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

    int   slot = -1, cur = TB_REF_DONE, pend = TB_REF_DONE;
    unsigned sp = bottom;                          // address of the next free entry
    int   best_rank = -1, best_tag = 0;
    float best_t = FLT_MAX;
    TbRay r;                                       // kx / ky / kz live packed in `kpack` between triangle rounds
    r.o = r.inv = v3s(0.f); r.sx = r.sy = r.sz = 0.f; r.kx = r.ky = r.kz = 0;
    int   kpack = 0;
    TbSelRay sr;
    sr.ix = sr.iy = sr.iz = sr.ox = sr.oy = sr.oz = 0.f;
    sr.nx = sr.ny = sr.nz = sr.fx = sr.fy = sr.fz = 0u;
    // bit 8 of kpack: the ray queue is empty — nothing to refill with, stay in the rounds to the end (a flag of its own
    // was one register too many: the compiler spilled it and re-read it every round)
    unsigned n_nodes = 0, n_tris = 0;
    unsigned n_rounds = 0, n_culled = 0, n_second = 0, n_parked = 0, n_empty = 0, n_tri_rounds = 0, n_refills = 0, n_refilled = 0;

    while (true) {
        // ---- refill idle lanes (warp converged here)
        const bool idle = cur == TB_REF_DONE && pend == TB_REF_DONE;
        const unsigned idle_mask = __ballot_sync(FULL, idle);
        if (idle_mask && !(kpack & 0x100)) {
            const int want = __popc(idle_mask);
            const int leader = __ffs(idle_mask) - 1;
            int base = 0;
            if ((int)lane == leader) base = atomicAdd(a.cursor, want);
            base = __shfl_sync(FULL, base, leader);
            if (base + want >= total) kpack |= 0x100;
            if (COUNT) ++n_refills;
            if (idle) {
                const int q = base + __popc(idle_mask & ((1u << lane) - 1u));
                if (q < total) {
                    slot = a.queue ? a.queue[q] : q;
                    const float4 o4 = tb_ld_stream(a.ray_o + slot), d4 = tb_ld_stream(a.ray_d + slot);
                    const V3 d = v3(d4.x, d4.y, d4.z);
                    r = tb_ray_setup(v3(o4.x, o4.y, o4.z), d);
                    kpack = (kpack & 0x100) | r.kx | (r.ky << 2) | (r.kz << 4);
                    if (!EXACT) sr = tb_sel_ray(r, d, a.grid_lo, a.grid_cell);
                    cur = 0; sp = bottom;
                    best_rank = -1; best_t = FLT_MAX;
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
            // 2. pop, resolved with selects. An entry the closest hit has culled since it was pushed leaves the lane
            //    in POP state: it pops again next round (a second pop in the same round, and a pop loop, were measured:
            //    -1.5 % and -10 %, experiments/README.md)
            {
                int2 e;
                if (SPILL && sp > smem_end) e = l_stack[(sp - smem_end) / ROW - 1];
                else e = tb_lds64(sp - ROW);
                const bool need = cur == TB_REF_POP;
                const int popped = __int_as_float(e.y) > cull_t ? TB_REF_POP : e.x;
                cur = need ? popped : cur;
                sp = need ? sp - ROW : sp;
            }
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
            float4 q0, q1, q2, q3, ta, tb, tc;
            uint4 qa, qb;
            if (is_node) {
                if (!EXACT) {
                    tb_ld_qnode(a.qnodes + (size_t)cur * 2, qa, qb);
                } else {
                    tb_load_node256(a.nodes + (size_t)cur * 4, q0, q1, q2, q3);
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
                if (!EXACT) {
                    ref0 = (int)qb.z; ref1 = (int)qb.w;
                    hit0 = tb_box_sel(sr, qa.x, qa.y, qa.z, &tmin0);
                    hit1 = tb_box_sel(sr, qa.w, qb.x, qb.y, &tmin1);
                } else {
                    ref0 = __float_as_int(q3.x); ref1 = __float_as_int(q3.y);
                    hit0 = tb_box(r, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, &tmin0);
                    hit1 = tb_box(r, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, &tmin1);
                }
                hit0 = hit0 && !(tmin0 > cull_t);
                hit1 = hit1 && !(tmin1 > cull_t);
                const bool near1 = tmin1 < tmin0;
                const bool both = hit0 && hit1;
                const int2 far = make_int2(near1 ? ref0 : ref1, __float_as_int(near1 ? tmin0 : tmin1));
                if (SPILL && sp >= smem_end) l_stack[(sp - smem_end) / ROW] = far;
                else tb_sts64(sp, far);                                    // scratch row when !both
                sp += both ? ROW : 0;
                const int one = hit0 ? ref0 : (hit1 ? ref1 : TB_REF_POP);
                cur = both ? (near1 ? ref1 : ref0) : one;
            }
            // 6. triangle step for the parked leaves
            if (do_tri) {
                const int rank = ~pend;
                if (COUNT) ++n_tris;
                float t, u, v;
                r.kx = kpack & 3; r.ky = (kpack >> 2) & 3; r.kz = (kpack >> 4) & 3;
                if (tb_triangle(r, v3(ta.x, ta.y, ta.z), v3(tb.x, tb.y, tb.z), v3(tc.x, tc.y, tc.z), &t, &u, &v)) {
                    if (t < best_t || (t == best_t && best_rank >= 0 && rank < best_rank)) {
                        best_t = t; best_rank = rank; best_tag = __float_as_int(tb.w);
                        // API rays: object | triangle << 8, u, v; wavefront: rank and the object | preset << 8 tag for the
                        // shade pass (u, v are not used there). A later, closer hit overwrites the record.
                        tb_st_stream(a.hit_out + slot, make_float4(t, a.emit_prim ? ta.w : __int_as_float(rank), a.emit_prim ? u : tb.w, v));
                    }
                }
                pend = TB_REF_DONE;
            }
            // 7. leave the rounds when the warp is done or has too few rays left to be worth a round
            const unsigned live = __ballot_sync(FULL, cur != TB_REF_DONE || pend != TB_REF_DONE);
            if (live == 0u || (!(kpack & 0x100) && __popc(live) < a.refetch_below)) break;
        }

        // ---- retire finished rays (warp converged)
        const bool finished = cur == TB_REF_DONE && pend == TB_REF_DONE && slot >= 0;
        if (__any_sync(FULL, finished)) {
            int preset = -1;
            if (finished) {
                if (best_rank >= 0) preset = (best_tag >> 8) & 0xff;
                else tb_st_stream(a.hit_out + slot, make_float4(-1.f, __int_as_float(-1), 0.f, 0.f));
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
