// experiments/tb_bvh_collapse4.cpp — NOT part of the build (see experiments/README.md): host side of the 4-wide
// tree of tb_trace_v5.cuh.
// ------------------------------------------------------------------------------------ 4-wide nodes
// tb_trace_v5.cuh's tree: every 4-wide node takes the two children of a binary node and, while it has
// fewer than four, replaces its inner child of largest surface area by that child's two children. Boxes go
// on the SAME 16-bit grid as the binary quantised nodes (so the kernels share the ray setup), with the same
// outward rounding. Leaves keep their ~rank refs. stack_bound = the deepest stack a traversal can need
// (each visited node leaves at most children-1 entries behind).
void tb_bvh_collapse4(const TbFlatBvh& flat, const TbQuantBvh& grid, TbQuantBvh4* out) {
    const size_t n2 = flat.nodes.size() / 16;
    struct Child { float lo[3], hi[3]; int32_t ref; };
    auto child_of = [&](int node, int c) {
        const float* q = &flat.nodes[(size_t)node * 16];
        Child ch;
        for (int k = 0; k < 3; ++k) { ch.lo[k] = q[6 * c + k]; ch.hi[k] = q[6 * c + 3 + k]; }
        memcpy(&ch.ref, &q[12 + c], 4);
        return ch;
    };
    auto area = [](const Child& c) {
        const double dx = (double)c.hi[0] - c.lo[0], dy = (double)c.hi[1] - c.lo[1], dz = (double)c.hi[2] - c.lo[2];
        return dx * dy + dy * dz + dz * dx;
    };
    out->nodes.clear();
    out->nodes.reserve(n2 / 2 * 16);
    std::vector<std::pair<int, int>> work;      // (binary node, stack entries left behind by the ancestors)
    work.push_back({ 0, 0 });
    out->stack_bound = 0;
    size_t next = 0;
    std::vector<int> src; src.push_back(0);      // binary source node of 4-wide node i
    std::vector<int> below; below.push_back(0);
    while (next < src.size()) {
        const int node = src[next], pending = below[next];
        Child ch[4];
        int cnt = 2;
        ch[0] = child_of(node, 0); ch[1] = child_of(node, 1);
        while (cnt < 4) {
            int pick = -1; double best = -1.0;
            for (int i = 0; i < cnt; ++i) if (ch[i].ref >= 0) { const double a_ = area(ch[i]); if (a_ > best) { best = a_; pick = i; } }
            if (pick < 0) break;
            const int inner = ch[pick].ref;
            ch[pick] = child_of(inner, 0);
            ch[cnt++] = child_of(inner, 1);
        }
        if (pending + cnt - 1 > out->stack_bound) out->stack_bound = pending + cnt - 1;
        uint32_t w[16];
        for (int c = 0; c < 4; ++c) {
            uint16_t g[6] = { 0, 0, 0, 0, 0, 0 };
            int32_t ref = (int32_t)0x80000000;          // TB_REF_DONE: empty slot
            if (c < cnt) {
                for (int k = 0; k < 3; ++k) {
                    const double l = grid.grid_lo[k], cell = grid.grid_cell[k];
                    long a = (long)floor((ch[c].lo[k] - l) / cell);
                    while (a > 0 && l + a * cell > ch[c].lo[k]) --a;
                    a -= 1;
                    long b = (long)ceil((ch[c].hi[k] - l) / cell);
                    while (b < 65535 && l + b * cell < ch[c].hi[k]) ++b;
                    b += 1;
                    if (a < 0) a = 0;
                    if (b > 65535) b = 65535;
                    g[k] = (uint16_t)a; g[3 + k] = (uint16_t)b;
                }
                ref = ch[c].ref;
                if (ref >= 0) {                          // inner child: becomes 4-wide node src.size()
                    const int id4 = (int)src.size();
                    src.push_back(ref);
                    below.push_back(pending + cnt - 1);
                    ref = id4;
                }
            }
            w[3 * c + 0] = (uint32_t)g[0] | ((uint32_t)g[1] << 16);
            w[3 * c + 1] = (uint32_t)g[2] | ((uint32_t)g[3] << 16);
            w[3 * c + 2] = (uint32_t)g[4] | ((uint32_t)g[5] << 16);
            memcpy(&w[12 + c], &ref, 4);
        }
        out->nodes.insert(out->nodes.end(), w, w + 16);
        ++next;
    }
    (void)n2;
}
