// tb_trace_packet.cuh — closest hit for COHERENT rays (camera rays): one warp walks the tree ONCE for its
// 32 rays.
//
// Why: the per-ray kernels spend 64 warp instructions per camera ray (ncu r1b, C2) although the 32 rays
// of a warp — an 8x4 pixel tile of one sample — visit nearly the same nodes: every lane keeps its own
// stack, fetches its own copy of the node and the warp diverges wherever two rays disagree. Here the
// warp shares ONE stack (shared memory, one entry per push), every node is fetched once at a
// warp-uniform address (one L1 wavefront), each lane tests the two child boxes against its own ray and
// own cull distance, and ballots decide where the packet goes: all control flow is warp-uniform. A ray
// pays for the union of the nodes its packet visits, divided by 32.
//
// Same answer as the per-ray kernels, by construction: a triangle is tested for a lane whenever the
// lane's ray enters the leaf's box within its cull distance (conservative culling only), the triangle
// test is the same tb_triangle, and the tie rule (equal t -> smaller rank) does not depend on the order
// of visits.
//   push: the far child goes on the stack with the MIN entry distance over the lanes that hit it;
//   pop:  an entry is skipped when that distance exceeds the MAX cull distance over the live lanes
//         (REDUX on the float bit patterns: all distances are >= 0, so they order as unsigned).
#pragma once

#define TB_PACKET_STACK (TB_STACK_SMEM + TB_STACK_LOCAL)

// GEN: the packet's rays are camera rays derived in registers from the slot index (tb_gen_ray) — nothing is read
// from the ray arrays, and the only path state written besides the hit is a zero radiance for the paths that leave
// the scene (the first shade pass initialises the others). Saves the 112 B per slot k_raygen wrote at the HBM
// write roof and the 32 B per slot this kernel read back (ncu r1w).
template <bool EXACT, bool GEN>
__global__ void __launch_bounds__(TB_TRACE_THREADS, TB_TRACE_CTAS_PER_SM) k_trace_packet(TbTraceArgs a, TbGenArgs g) {
    __shared__ int   s_ref[TB_TRACE_THREADS / 32][TB_PACKET_STACK];
    __shared__ float s_tmin[TB_TRACE_THREADS / 32][TB_PACKET_STACK];
    const int tid = threadIdx.x;
    const unsigned lane = tid & 31u;
    const int warp = tid >> 5;
    const unsigned FULL = 0xffffffffu;
    const int total = a.count_ptr ? *a.count_ptr : a.count;
    int*   const st_ref = s_ref[warp];
    float* const st_tmin = s_tmin[warp];

    while (true) {
        int base = 0;
        if (lane == 0) base = atomicAdd(a.cursor, 32);
        base = __shfl_sync(FULL, base, 0);
        if (base >= total) break;
        const int q = base + (int)lane;
        const bool alive = q < total;
        int slot = -1;
        TbRay r;
        r.o = r.inv = v3s(0.f); r.sx = r.sy = r.sz = 0.f; r.kx = r.ky = r.kz = 0;
        TbFastRay fr;
        fr.ix = fr.iy = fr.iz = fr.ox = fr.oy = fr.oz = 0.f;
        if (alive) {
            V3 o, d;
            if (GEN) {
                slot = q;
                uint64_t key;
                d = tb_gen_ray(g, slot, &key, &o);
            } else {
                slot = a.queue ? a.queue[q] : q;
                const float4 o4 = a.ray_o[slot], d4 = a.ray_d[slot];
                o = v3(o4.x, o4.y, o4.z); d = v3(d4.x, d4.y, d4.z);
            }
            r = tb_ray_setup(o, d);
            if (!EXACT) fr = tb_fast_ray(r, d);
        }
        int   best_rank = -1;
        float best_t = FLT_MAX, best_u = 0.f, best_v = 0.f;
        float cull_t = alive ? FLT_MAX : 0.f;             // dead lanes must not hold the warp's max cull distance
        int   sp = 0;
        int   node = 0;

#define TB_PACKET_LEAF(REF, HIT)                                                                    \
        do {                                                                                        \
            const int rank_ = ~(REF);                                                               \
            const float4* tp_ = a.tris + (size_t)rank_ * 3;                                         \
            const float4 ta_ = __ldg(tp_), tb_ = __ldg(tp_ + 1), tc_ = __ldg(tp_ + 2);              \
            float t_, u_, v_;                                                                       \
            if ((HIT) && tb_triangle(r, v3(ta_.x, ta_.y, ta_.z), v3(tb_.x, tb_.y, tb_.z), v3(tc_.x, tc_.y, tc_.z), &t_, &u_, &v_)) { \
                if (t_ < best_t || (t_ == best_t && best_rank >= 0 && rank_ < best_rank)) {         \
                    best_t = t_; best_rank = rank_; best_u = u_; best_v = v_;                       \
                    cull_t = best_t + best_t * 1e-5f;                                               \
                }                                                                                   \
            }                                                                                       \
        } while (0)

        while (true) {
            // ---- one inner node for the whole packet (uniform address: one wavefront per load)
            const float4* np = a.nodes + (size_t)node * 4;
            const float4 q0 = __ldg(np), q1 = __ldg(np + 1), q2 = __ldg(np + 2), q3 = __ldg(np + 3);
            const int ref0 = __float_as_int(q3.x), ref1 = __float_as_int(q3.y);
            float tmin0, tmin1;
            bool hit0, hit1;
            if (EXACT) {
                hit0 = tb_box(r, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, &tmin0);
                hit1 = tb_box(r, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, &tmin1);
            } else {
                hit0 = tb_box_fast(fr, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, &tmin0);
                hit1 = tb_box_fast(fr, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, &tmin1);
            }
            hit0 = hit0 && alive && !(tmin0 > cull_t);
            hit1 = hit1 && alive && !(tmin1 > cull_t);
            unsigned m0 = __ballot_sync(FULL, hit0), m1 = __ballot_sync(FULL, hit1);
            // ---- leaf children are tested at once (nearer one first when both are leaves: the first
            //      hit then culls the second for most lanes)
            bool tested = false;
            if (ref0 < 0 && ref1 < 0 && m0 && m1) {
                const bool one_first = __popc(__ballot_sync(FULL, hit1 && (!hit0 || tmin1 < tmin0))) > __popc(m0) / 2;
                if (one_first) { TB_PACKET_LEAF(ref1, hit1); hit0 = hit0 && !(tmin0 > cull_t); TB_PACKET_LEAF(ref0, hit0); }
                else           { TB_PACKET_LEAF(ref0, hit0); hit1 = hit1 && !(tmin1 > cull_t); TB_PACKET_LEAF(ref1, hit1); }
                m0 = m1 = 0u;
            } else {
                if (ref0 < 0 && m0) { TB_PACKET_LEAF(ref0, hit0); m0 = 0u; tested = true; }
                if (ref1 < 0 && m1) { TB_PACKET_LEAF(ref1, hit1); m1 = 0u; tested = true; }
                if (tested) {                              // the hit just found may cull the inner sibling
                    if (m0) { hit0 = hit0 && !(tmin0 > cull_t); m0 = __ballot_sync(FULL, hit0); }
                    if (m1) { hit1 = hit1 && !(tmin1 > cull_t); m1 = __ballot_sync(FULL, hit1); }
                }
            }
            // ---- where does the packet go?
            if (m0 && m1) {
                const unsigned want1 = __ballot_sync(FULL, hit1 && (!hit0 || tmin1 < tmin0));
                const unsigned want0 = __ballot_sync(FULL, hit0 && (!hit1 || !(tmin1 < tmin0)));
                const bool near1 = __popc(want1) > __popc(want0);
                const int  far_ref = near1 ? ref0 : ref1;
                const bool far_hit = near1 ? hit0 : hit1;
                const float far_t = fmaxf(near1 ? tmin0 : tmin1, 0.f);   // EXACT boxes report negative / NaN entry distances as they are
                const unsigned far_min = __reduce_min_sync(FULL, far_hit ? __float_as_uint(far_t) : 0x7f800000u);
                if (lane == 0) { st_ref[sp] = far_ref; st_tmin[sp] = __uint_as_float(far_min); }
                __syncwarp();
                ++sp;
                node = near1 ? ref1 : ref0;
            } else if (m0) {
                node = ref0;
            } else if (m1) {
                node = ref1;
            } else {
                const unsigned cull_max = __reduce_max_sync(FULL, __float_as_uint(cull_t));
                node = -1;
                while (sp > 0) {
                    --sp;
                    const int   n_ = st_ref[sp];
                    const float tm_ = st_tmin[sp];
                    if (!(__float_as_uint(tm_) > cull_max)) { node = n_; break; }
                }
                if (node < 0) break;
            }
        }
#undef TB_PACKET_LEAF

        // ---- retire the packet
        int preset = -1;
        if (alive) {
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
                if (GEN) a.lo[slot] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            a.hit_out[slot] = h;
        }
        if (a.matq) {
#pragma unroll
            for (int p = 0; p < TB_PRESET_COUNT; ++p) {
                const unsigned m = __ballot_sync(FULL, alive && preset == p);
                if (m) {
                    const int leader = __ffs(m) - 1;
                    int b0 = 0;
                    if ((int)lane == leader) b0 = atomicAdd(a.matq_count + p, __popc(m));
                    b0 = __shfl_sync(FULL, b0, leader);
                    if (alive && preset == p) a.matq[(size_t)p * a.matq_stride + b0 + __popc(m & ((1u << lane) - 1u))] = slot;
                }
            }
        }
        __syncwarp();
    }
}
