/* terra_headless — headless front-end for the Terra C API.
 *
 * The reference's only front-end is Satellite, a Windows GL/ImGui viewer (satellite/src): it imports an
 * OBJ/MTL model (satellite/include/Apollo.h), maps every mesh to one TerraObject with a Diffuse or Phong
 * material (satellite/src/Scene.cpp:144-237), commits, renders tile by tile (Renderer.cpp:316-349) and
 * saves the framebuffer (Visualization.cpp:286-349). This is the same caller without the window: plain C11
 * against the Terra API only (no terra_b200_* extension), so ONE source builds against libterra_b200.so
 * (GPU) and against the unmodified reference library (CPU) — the drop-in property as a program.
 *
 *   terra_headless [--obj model.obj] [--width W --height H] [--spp N] [--bounces N]
 *                  [--integrator simple|direct|mis|mono|depth|normals|misweights] [--jitter J]
 *                  [--camera px py pz dx dy dz] [--fov deg] [--tonemap none|linear|reinhard|filmic|uncharted2]
 *                  [--exposure E] [--gamma G] [--env r g b] [--tile N] [--passes N] [--progress]
 *                  [--config satellite.config] [--out image.pfm|.ppm|.png|.hdr] [--dry-run]
 *                  [--scene cornell|field] [--devices 0,1,...  (product build only: multi-GPU fan-out behind terra_render)]
 * --config reads the reference front-end's `name = value` file (satellite/include/Config.hpp:19-117,
 * satellite/src/Config.cpp:170-231); options are applied in command-line order. .png / .hdr are written the way
 * satellite's save_to_file does (Visualization.cpp:286-349): 8-bit clamp(v)*255 and Radiance RGBE of pixels[].
 * Without --obj a built-in Cornell box is rendered. --tile N renders the frame as NxN-pixel terra_render
 * calls (the reference caller's pattern); the default is one full-frame call. --dry-run builds the scene
 * and prints what it would render, without terra_scene_commit (needs no GPU).
 */
#include <ctype.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../include/terra_b200.h"

/* ------------------------------------------------------------------ growable arrays */
typedef struct { float* v; size_t n, cap; } FloatVec;
static void fpush(FloatVec* a, float x) { if (a->n == a->cap) { a->cap = a->cap ? a->cap * 2 : 1024; a->v = realloc(a->v, a->cap * sizeof(float)); } a->v[a->n++] = x; }

/* ------------------------------------------------------------------ MTL */
typedef struct { char name[128]; float kd[3], ks[3], ke[3], ns; int illum; } Mtl;
typedef struct { Mtl* v; size_t n, cap; } MtlVec;

static Mtl* mtl_find(MtlVec* m, const char* name) {
    for (size_t i = 0; i < m->n; ++i) if (strcmp(m->v[i].name, name) == 0) return &m->v[i];
    return NULL;
}
static Mtl* mtl_add(MtlVec* m, const char* name) {
    if (m->n == m->cap) { m->cap = m->cap ? m->cap * 2 : 16; m->v = realloc(m->v, m->cap * sizeof(Mtl)); }
    Mtl* x = &m->v[m->n++];
    memset(x, 0, sizeof *x);
    snprintf(x->name, sizeof x->name, "%s", name);
    x->kd[0] = x->kd[1] = x->kd[2] = 0.7f; x->ns = 10.f; x->illum = 1;
    return x;
}
static void load_mtl(const char* path, MtlVec* out) {
    FILE* f = fopen(path, "r");
    if (!f) { fprintf(stderr, "terra_headless: cannot open material library %s (materials default to grey diffuse)\n", path); return; }
    char line[1024]; Mtl* cur = NULL;
    while (fgets(line, sizeof line, f)) {
        char key[64], name[512];
        if (sscanf(line, "%63s", key) != 1 || key[0] == '#') continue;
        if (strcmp(key, "newmtl") == 0 && sscanf(line, "%*s %511s", name) == 1) cur = mtl_add(out, name);
        else if (!cur) continue;
        else if (strcmp(key, "Kd") == 0) sscanf(line, "%*s %f %f %f", &cur->kd[0], &cur->kd[1], &cur->kd[2]);
        else if (strcmp(key, "Ks") == 0) sscanf(line, "%*s %f %f %f", &cur->ks[0], &cur->ks[1], &cur->ks[2]);
        else if (strcmp(key, "Ke") == 0) sscanf(line, "%*s %f %f %f", &cur->ke[0], &cur->ke[1], &cur->ke[2]);
        else if (strcmp(key, "Ns") == 0) sscanf(line, "%*s %f", &cur->ns);
        else if (strcmp(key, "illum") == 0) sscanf(line, "%*s %d", &cur->illum);
    }
    fclose(f);
}

/* ------------------------------------------------------------------ OBJ -> one triangle list per material */
typedef struct { int mtl; FloatVec pos, nrm, tex; } Group;          /* 9 + 9 + 6 floats per triangle */
typedef struct { Group* v; size_t n, cap; } GroupVec;

static Group* group_for(GroupVec* g, int mtl) {
    for (size_t i = 0; i < g->n; ++i) if (g->v[i].mtl == mtl) return &g->v[i];
    if (g->n == g->cap) { g->cap = g->cap ? g->cap * 2 : 8; g->v = realloc(g->v, g->cap * sizeof(Group)); }
    Group* x = &g->v[g->n++];
    memset(x, 0, sizeof *x);
    x->mtl = mtl;
    return x;
}
static int obj_index(int i, size_t count) { return i > 0 ? i - 1 : (i < 0 ? (int)count + i : -1); }

static int load_obj(const char* path, GroupVec* groups, MtlVec* mtls) {
    FILE* f = fopen(path, "r");
    if (!f) { fprintf(stderr, "terra_headless: cannot open %s\n", path); return 0; }
    char dir[1024]; snprintf(dir, sizeof dir, "%s", path);
    char* slash = strrchr(dir, '/'); if (slash) slash[1] = 0; else dir[0] = 0;
    FloatVec P = {0}, N = {0}, T = {0};
    int cur_mtl = -1;
    char* line = malloc(1 << 16);
    while (fgets(line, 1 << 16, f)) {
        char* s = line; while (isspace((unsigned char)*s)) ++s;
        if (s[0] == 'v' && s[1] == ' ') { float x, y, z; if (sscanf(s + 2, "%f %f %f", &x, &y, &z) == 3) { fpush(&P, x); fpush(&P, y); fpush(&P, z); } }
        else if (s[0] == 'v' && s[1] == 'n') { float x, y, z; if (sscanf(s + 3, "%f %f %f", &x, &y, &z) == 3) { fpush(&N, x); fpush(&N, y); fpush(&N, z); } }
        else if (s[0] == 'v' && s[1] == 't') { float u = 0, v = 0; sscanf(s + 3, "%f %f", &u, &v); fpush(&T, u); fpush(&T, v); }
        else if (strncmp(s, "mtllib", 6) == 0) { char name[512], full[1600]; if (sscanf(s + 6, "%511s", name) == 1) { snprintf(full, sizeof full, "%s%s", dir, name); load_mtl(full, mtls); } }
        else if (strncmp(s, "usemtl", 6) == 0) {
            char name[512];
            if (sscanf(s + 6, "%511s", name) == 1) { Mtl* m = mtl_find(mtls, name); if (!m) m = mtl_add(mtls, name); cur_mtl = (int)(m - mtls->v); }
        } else if (s[0] == 'f' && isspace((unsigned char)s[1])) {
            int vi[64], ti[64], ni[64], n = 0;
            char* tok = strtok(s + 2, " \t\r\n");
            while (tok && n < 64) {
                int a = 0, b = 0, c = 0;
                if (sscanf(tok, "%d/%d/%d", &a, &b, &c) == 3) {}
                else if (sscanf(tok, "%d//%d", &a, &c) == 2) { b = 0; }
                else if (sscanf(tok, "%d/%d", &a, &b) == 2) { c = 0; }
                else if (sscanf(tok, "%d", &a) == 1) { b = c = 0; }
                else break;
                vi[n] = obj_index(a, P.n / 3); ti[n] = obj_index(b, T.n / 2); ni[n] = obj_index(c, N.n / 3);
                ++n;
                tok = strtok(NULL, " \t\r\n");
            }
            Group* g = group_for(groups, cur_mtl);
            for (int k = 1; k + 1 < n; ++k) {                     /* triangle fan */
                const int c3[3] = { 0, k, k + 1 };
                float p[3][3], fn[3];
                for (int j = 0; j < 3; ++j) {
                    const int v = vi[c3[j]];
                    if (v < 0 || (size_t)v >= P.n / 3) { fprintf(stderr, "terra_headless: face index out of range in %s\n", path); fclose(f); return 0; }
                    for (int d = 0; d < 3; ++d) p[j][d] = P.v[3 * v + d];
                }
                { const float e0[3] = { p[1][0]-p[0][0], p[1][1]-p[0][1], p[1][2]-p[0][2] }, e1[3] = { p[2][0]-p[0][0], p[2][1]-p[0][1], p[2][2]-p[0][2] };
                  fn[0] = e0[1]*e1[2]-e0[2]*e1[1]; fn[1] = e0[2]*e1[0]-e0[0]*e1[2]; fn[2] = e0[0]*e1[1]-e0[1]*e1[0];
                  float l = sqrtf(fn[0]*fn[0]+fn[1]*fn[1]+fn[2]*fn[2]); if (l > 0) { fn[0] /= l; fn[1] /= l; fn[2] /= l; } }
                for (int j = 0; j < 3; ++j) {
                    for (int d = 0; d < 3; ++d) fpush(&g->pos, p[j][d]);
                    const int nn = ni[c3[j]], tt = ti[c3[j]];
                    for (int d = 0; d < 3; ++d) fpush(&g->nrm, (nn >= 0 && (size_t)nn < N.n / 3) ? N.v[3 * nn + d] : fn[d]);
                    for (int d = 0; d < 2; ++d) fpush(&g->tex, (tt >= 0 && (size_t)tt < T.n / 2) ? T.v[2 * tt + d] : 0.f);
                }
            }
        }
    }
    free(line); free(P.v); free(N.v); free(T.v);
    fclose(f);
    return 1;
}

/* Built-in Cornell box as the same group structure (unit box, emissive ceiling quad). */
static void add_quad(Group* g, const float a[3], const float b[3], const float c[3], const float d[3]) {
    const float* tri[2][3] = { { a, b, c }, { a, c, d } };
    const float e0[3] = { b[0]-a[0], b[1]-a[1], b[2]-a[2] }, e1[3] = { c[0]-a[0], c[1]-a[1], c[2]-a[2] };
    float n[3] = { e0[1]*e1[2]-e0[2]*e1[1], e0[2]*e1[0]-e0[0]*e1[2], e0[0]*e1[1]-e0[1]*e1[0] };
    float l = sqrtf(n[0]*n[0]+n[1]*n[1]+n[2]*n[2]); n[0] /= l; n[1] /= l; n[2] /= l;
    for (int t = 0; t < 2; ++t) for (int j = 0; j < 3; ++j) {
        for (int k = 0; k < 3; ++k) { fpush(&g->pos, tri[t][j][k]); fpush(&g->nrm, n[k]); }
        fpush(&g->tex, 0.f); fpush(&g->tex, 0.f);
    }
}
static void builtin_cornell(GroupVec* groups, MtlVec* mtls) {
    const float W[3] = { .73f, .73f, .73f }, R[3] = { .65f, .05f, .05f }, G[3] = { .12f, .45f, .15f };
    Mtl* m;
    m = mtl_add(mtls, "white"); memcpy(m->kd, W, 12);
    m = mtl_add(mtls, "red");   memcpy(m->kd, R, 12);
    m = mtl_add(mtls, "green"); memcpy(m->kd, G, 12);
    m = mtl_add(mtls, "light"); m->kd[0] = m->kd[1] = m->kd[2] = 0.f; m->ke[0] = m->ke[1] = m->ke[2] = 15.f;
    const float p000[3] = {0,0,0}, p100[3] = {1,0,0}, p010[3] = {0,1,0}, p110[3] = {1,1,0}, p001[3] = {0,0,1}, p101[3] = {1,0,1}, p011[3] = {0,1,1}, p111[3] = {1,1,1};
    Group* w = group_for(groups, 0);
    /* windings chosen so the geometric normals face INTO the box (the presets sample around the shading normal) */
    add_quad(w, p000, p001, p101, p100);            /* floor   +y */
    add_quad(w, p010, p110, p111, p011);            /* ceiling -y */
    add_quad(w, p001, p011, p111, p101);            /* back    -z */
    add_quad(group_for(groups, 1), p000, p010, p011, p001);   /* left  +x: red */
    add_quad(group_for(groups, 2), p100, p101, p111, p110);   /* right -x: green */
    const float l0[3] = {.35f,.999f,.35f}, l1[3] = {.65f,.999f,.35f}, l2[3] = {.65f,.999f,.65f}, l3[3] = {.35f,.999f,.65f};
    add_quad(group_for(groups, 3), l0, l1, l2, l3);
}

/* --scene field: a procedural field of uv-spheres on a jittered grid over a floor (the shape of BASELINE config C2:
 * 125 spheres x 760 triangles = 95k, 16 Lambert groups, 4 emissive spheres), so the multi-GPU fan-out has a workload
 * without an asset file. Plain C generator (own LCG): same class of scene as terra_b200/scenes.py, not its bytes. */
static unsigned field_rng = 12345u;
static float field_rand(void) { field_rng = field_rng * 1664525u + 1013904223u; return (float)(field_rng >> 8) / 16777216.f; }
static void add_tri_n(Group* g, const float* a, const float* b, const float* c, const float* ctr, float r) {
    const float* p[3] = { a, b, c };
    for (int k = 0; k < 3; ++k) for (int j = 0; j < 3; ++j) fpush(&g->pos, p[k][j]);
    for (int k = 0; k < 3; ++k) for (int j = 0; j < 3; ++j) fpush(&g->nrm, (p[k][j] - ctr[j]) / r);
    for (int k = 0; k < 6; ++k) fpush(&g->tex, 0.f);
}
static void builtin_field(GroupVec* groups, MtlVec* mtls, float* cam) {
    const float extent = 100.f; const int gdim = 5, n_lat = 20, n_lon = 20, n_groups = 16;
    for (int i = 0; i < n_groups; ++i) { char nm[32]; snprintf(nm, sizeof nm, "lambert%d", i); Mtl* m = mtl_add(mtls, nm); for (int k = 0; k < 3; ++k) m->kd[k] = 0.2f + 0.6f * field_rand(); }
    { Mtl* m = mtl_add(mtls, "emitter"); m->ke[0] = 12.f; m->ke[1] = 11.f; m->ke[2] = 10.f; }
    { Mtl* m = mtl_add(mtls, "floor"); m->kd[0] = m->kd[1] = m->kd[2] = 0.5f; }
    const float cell = extent / gdim, PI = 3.14159265358979f;
    int idx = 0, emitters = 0;
    for (int i = 0; i < gdim; ++i) for (int j = 0; j < gdim; ++j) for (int k = 0; k < gdim; ++k, ++idx) {
        const float ctr[3] = { (i + 0.3f + 0.4f * field_rand()) * cell, (j + 0.3f + 0.4f * field_rand()) * cell, (k + 0.3f + 0.4f * field_rand()) * cell };
        const float r = cell * (0.22f + 0.10f * field_rand());
        int grp = idx % n_groups;
        if (idx % 31 == 7 && emitters < 4) { grp = n_groups; ++emitters; }
        Group* g = group_for(groups, grp);
        for (int a = 0; a < n_lat; ++a) for (int b = 0; b < n_lon; ++b) {
            float v[4][3];
            for (int q = 0; q < 4; ++q) {
                const float th = PI * (a + (q >> 1)) / n_lat, ph = 2 * PI * (b + (q & 1)) / n_lon;
                v[q][0] = ctr[0] + r * sinf(th) * cosf(ph); v[q][1] = ctr[1] + r * cosf(th); v[q][2] = ctr[2] + r * sinf(th) * sinf(ph);
            }
            if (a != 0) add_tri_n(g, v[0], v[2], v[1], ctr, r);
            if (a != n_lat - 1) add_tri_n(g, v[1], v[2], v[3], ctr, r);
        }
    }
    const float m = 0.15f * extent, y = -0.5f;
    const float f0[3] = { -m, y, -m }, f1[3] = { extent + m, y, -m }, f2[3] = { extent + m, y, extent + m }, f3[3] = { -m, y, extent + m };
    add_quad(group_for(groups, n_groups + 1), f0, f3, f2, f1);
    const float c[6] = { extent * 0.5f, extent * 0.62f, -extent * 1.05f, 0.f, -0.18f, 1.f };
    memcpy(cam, c, sizeof c);
}

/* ------------------------------------------------------------------ image output */
/* PNG without a compression library: 8-bit RGB, filter 0, zlib stream of STORED deflate blocks.
 * Bytes as satellite's save_to_file makes them (Visualization.cpp:325-333): (uint8_t)(clamp(v, 0, 1) * 255.f). */
static unsigned long crc_table[256];
static void crc_init(void) {
    for (unsigned long n = 0; n < 256; ++n) { unsigned long c = n; for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xedb88320UL ^ (c >> 1) : c >> 1; crc_table[n] = c; }
}
static unsigned long crc_update(unsigned long c, const unsigned char* b, size_t n) {
    for (size_t i = 0; i < n; ++i) c = crc_table[(c ^ b[i]) & 0xff] ^ (c >> 8);
    return c;
}
static void be32(unsigned char* p, unsigned long v) { p[0] = (unsigned char)(v >> 24); p[1] = (unsigned char)(v >> 16); p[2] = (unsigned char)(v >> 8); p[3] = (unsigned char)v; }
static void png_chunk(FILE* f, const char* tag, const unsigned char* data, size_t n) {
    unsigned char h[8];
    be32(h, (unsigned long)n); memcpy(h + 4, tag, 4);
    fwrite(h, 8, 1, f);
    if (n) fwrite(data, n, 1, f);
    unsigned long c = crc_update(0xffffffffUL, h + 4, 4);
    c = crc_update(c, data, n) ^ 0xffffffffUL;
    be32(h, c);
    fwrite(h, 4, 1, f);
}
static int save_png(FILE* f, const TerraFramebuffer* fb) {
    const size_t w = fb->width, h = fb->height, row = 1 + 3 * w, raw_n = row * h;
    unsigned char* raw = (unsigned char*)malloc(raw_n);
    const size_t blocks = (raw_n + 65534) / 65535;
    unsigned char* z = (unsigned char*)malloc(2 + raw_n + 5 * blocks + 4);
    if (!raw || !z) return 0;
    for (size_t y = 0; y < h; ++y) {
        raw[y * row] = 0;
        for (size_t x = 0; x < w; ++x) {
            const float c[3] = { fb->pixels[y * w + x].x, fb->pixels[y * w + x].y, fb->pixels[y * w + x].z };
            for (int k = 0; k < 3; ++k) { float v = c[k] > 0.f ? c[k] : 0.f; v = v < 1.f ? v : 1.f; raw[y * row + 1 + 3 * x + k] = (unsigned char)(v * 255.f); }
        }
    }
    size_t zn = 0;
    z[zn++] = 0x78; z[zn++] = 0x01;
    unsigned long a = 1, b = 0;
    for (size_t at = 0; at < raw_n; at += 65535) {
        const size_t n = raw_n - at < 65535 ? raw_n - at : 65535;
        z[zn++] = at + n == raw_n ? 1 : 0;
        z[zn++] = (unsigned char)(n & 0xff); z[zn++] = (unsigned char)(n >> 8);
        z[zn++] = (unsigned char)(~n & 0xff); z[zn++] = (unsigned char)((~n >> 8) & 0xff);
        memcpy(z + zn, raw + at, n); zn += n;
        for (size_t i = 0; i < n; ++i) { a = (a + raw[at + i]) % 65521; b = (b + a) % 65521; }
    }
    be32(z + zn, (b << 16) | a); zn += 4;
    static const unsigned char sig[8] = { 0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a };
    unsigned char ihdr[13];
    crc_init();
    fwrite(sig, 8, 1, f);
    be32(ihdr, (unsigned long)w); be32(ihdr + 4, (unsigned long)h);
    ihdr[8] = 8; ihdr[9] = 2; ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0;
    png_chunk(f, "IHDR", ihdr, 13);
    png_chunk(f, "IDAT", z, zn);
    png_chunk(f, "IEND", NULL, 0);
    free(raw); free(z);
    return 1;
}
/* Radiance .hdr (what stbi_write_hdr gives satellite, Visualization.cpp:341): RGBE, flat scanlines, top row first. */
static int save_hdr(FILE* f, const TerraFramebuffer* fb) {
    fprintf(f, "#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %zu +X %zu\n", fb->height, fb->width);
    for (size_t i = 0; i < fb->width * fb->height; ++i) {
        const float c[3] = { fb->pixels[i].x, fb->pixels[i].y, fb->pixels[i].z };
        const float m = fmaxf(c[0], fmaxf(c[1], c[2]));
        unsigned char rgbe[4] = { 0, 0, 0, 0 };
        if (m >= 1e-32f) {
            int e;
            const float scale = frexpf(m, &e) * 256.f / m;
            for (int k = 0; k < 3; ++k) rgbe[k] = (unsigned char)(c[k] > 0.f ? c[k] * scale : 0.f);
            rgbe[3] = (unsigned char)(e + 128);
        }
        fwrite(rgbe, 4, 1, f);
    }
    return 1;
}
static int save_image(const char* path, const TerraFramebuffer* fb, int hdr_radiance) {
    FILE* f = fopen(path, "wb");
    if (!f) { fprintf(stderr, "terra_headless: cannot write %s\n", path); return 0; }
    const size_t w = fb->width, h = fb->height;
    const char* ext = strrchr(path, '.');
    if (ext && strcmp(ext, ".png") == 0) { const int ok = save_png(f, fb); fclose(f); return ok; }
    if (ext && strcmp(ext, ".hdr") == 0) { const int ok = save_hdr(f, fb); fclose(f); return ok; }
    if (ext && strcmp(ext, ".pfm") == 0) {                  /* float RGB, bottom row first, little endian */
        fprintf(f, "PF\n%zu %zu\n-1.0\n", w, h);
        for (size_t y = h; y-- > 0;)
            for (size_t x = 0; x < w; ++x) {
                float rgb[3];
                if (hdr_radiance) {                          /* acc / samples: what Visualization saves as HDR */
                    const TerraRawIntegrationResult* r = &fb->results[y * w + x];
                    const float inv = r->samples > 0 ? 1.f / (float)r->samples : 0.f;
                    rgb[0] = r->acc.x * inv; rgb[1] = r->acc.y * inv; rgb[2] = r->acc.z * inv;
                } else { rgb[0] = fb->pixels[y * w + x].x; rgb[1] = fb->pixels[y * w + x].y; rgb[2] = fb->pixels[y * w + x].z; }
                fwrite(rgb, sizeof rgb, 1, f);
            }
    } else {                                                /* 8-bit PPM of the tonemapped pixels[] */
        fprintf(f, "P6\n%zu %zu\n255\n", w, h);
        for (size_t i = 0; i < w * h; ++i) {
            const float c[3] = { fb->pixels[i].x, fb->pixels[i].y, fb->pixels[i].z };
            unsigned char b[3];
            for (int k = 0; k < 3; ++k) { float v = c[k] < 0 ? 0 : (c[k] > 1 ? 1 : c[k]); b[k] = (unsigned char)(v * 255.f + 0.5f); }
            fwrite(b, 3, 1, f);
        }
    }
    fclose(f);
    return 1;
}

static int pick(const char* v, const char* const* names, int n, const char* what) {
    for (int i = 0; i < n; ++i) if (strcmp(v, names[i]) == 0) return i;
    fprintf(stderr, "terra_headless: unknown %s '%s'\n", what, v);
    exit(2);
}

static const char* const integ[] = { "simple", "direct", "mis", "mono", "depth", "normals", "misweights" };
static const char* const tones[] = { "none", "linear", "reinhard", "filmic", "uncharted2" };

/* ------------------------------------------------------------------ satellite.config
 * `name = value` lines, '#' comments, blank lines skipped (satellite/src/Config.cpp:170-231); names and value
 * spellings of satellite/include/Config.hpp:19-117. workers / progressive / accelerator are accepted and have
 * nothing to drive here. Returns 0 on a malformed value (the reference logs and keeps the default). */
typedef struct {
    char obj[1024]; int have_obj;
    size_t width, height, spp, bounces, tile;
    int integrator, tonemap, sampler, have_cam, have_up;
    float jitter, fov, exposure, gamma_, env[3], cam[6], up[3];
} Settings;

static int parse_f3(const char* v, float* out) {
    char* end = NULL;
    out[0] = strtof(v, &end); out[1] = strtof(end, &end); out[2] = strtof(end, &end);
    while (*end == ' ' || *end == '\t' || *end == '\r') ++end;
    return *end == '\0';
}

static int config_set(Settings* S, const char* name, const char* value) {
    char* end = NULL;
    if (!strcmp(name, "scene")) {
        snprintf(S->obj, sizeof S->obj, "%s", value);
        for (char* c = S->obj; *c; ++c) if (*c == '\\') *c = '/';          /* the shipped config uses Windows separators */
        S->have_obj = 1; return 1;
    }
    if (!strcmp(name, "tile-size")) { const size_t v = strtoul(value, &end, 10); if (*value == '\0' || *end != '\0') return 0; S->tile = v; return 1; }
    if (!strcmp(name, "bounces")) { const size_t v = strtoul(value, &end, 10); if (*value == '\0' || *end != '\0') return 0; S->bounces = v; return 1; }
    if (!strcmp(name, "samples")) { const size_t v = strtoul(value, &end, 10); if (*value == '\0' || *end != '\0') return 0; S->spp = v; return 1; }
    if (!strcmp(name, "width")) { const size_t v = strtoul(value, &end, 10); if (*value == '\0' || *end != '\0') return 0; S->width = v; return 1; }
    if (!strcmp(name, "height")) { const size_t v = strtoul(value, &end, 10); if (*value == '\0' || *end != '\0') return 0; S->height = v; return 1; }
    if (!strcmp(name, "gamma")) { const float v = strtof(value, &end); if (*value == '\0' || *end != '\0') return 0; S->gamma_ = v; return 1; }
    if (!strcmp(name, "exposure")) { const float v = strtof(value, &end); if (*value == '\0' || *end != '\0') return 0; S->exposure = v; return 1; }
    if (!strcmp(name, "jitter")) { const float v = strtof(value, &end); if (*value == '\0' || *end != '\0') return 0; S->jitter = v; return 1; }
    if (!strcmp(name, "camfov")) { const float v = strtof(value, &end); if (*value == '\0' || *end != '\0') return 0; S->fov = v; return 1; }
    float v3[3];
    if (!strcmp(name, "campos")) { if (!parse_f3(value, v3)) return 0; memcpy(S->cam, v3, sizeof v3); S->have_cam = 1; return 1; }
    if (!strcmp(name, "camdir")) { if (!parse_f3(value, v3)) return 0; memcpy(S->cam + 3, v3, sizeof v3); S->have_cam = 1; return 1; }
    if (!strcmp(name, "camup")) { if (!parse_f3(value, v3)) return 0; memcpy(S->up, v3, sizeof v3); S->have_up = 1; return 1; }
    if (!strcmp(name, "envmap")) { if (!parse_f3(value, v3)) return 0; memcpy(S->env, v3, sizeof v3); return 1; }
    if (!strcmp(name, "tonemap")) {
        static const char* const names[] = { "none", "linear", "reinhard", "filmic", "uncharted" };
        for (int i = 0; i < 5; ++i) if (!strcmp(value, names[i])) { S->tonemap = i; return 1; }
        return 0;
    }
    if (!strcmp(name, "integrator")) {
        static const char* const names[] = { "simple", "direct", "mis", "debug-mono", "debug-depth", "debug-normals", "debug-mis" };
        for (int i = 0; i < 7; ++i) if (!strcmp(value, names[i])) { S->integrator = i; return 1; }
        return 0;
    }
    if (!strcmp(name, "sampler")) {
        static const char* const names[] = { "random", "stratified", "halton" };
        for (int i = 0; i < 3; ++i) if (!strcmp(value, names[i])) { S->sampler = i; return 1; }
        return 0;
    }
    if (!strcmp(name, "workers") || !strcmp(name, "progressive") || !strcmp(name, "accelerator")) return 1;
    fprintf(stderr, "terra_headless: config: unknown option '%s'\n", name);
    return 1;
}

static int load_config(Settings* S, const char* path) {
    FILE* f = fopen(path, "r");
    if (!f) { fprintf(stderr, "terra_headless: cannot read %s\n", path); return 0; }
    char line[2048];
    while (fgets(line, sizeof line, f)) {
        char* p = line;
        while (*p == ' ' || *p == '\t') ++p;
        if (*p == '#' || *p == '\n' || *p == '\r' || *p == '\0') continue;
        char* eq = strchr(p, '=');
        if (!eq) { fprintf(stderr, "terra_headless: config: unrecognised name=value line: %s", line); continue; }
        char* name_end = eq;
        while (name_end > p && (name_end[-1] == ' ' || name_end[-1] == '\t')) --name_end;
        *name_end = '\0';
        char* v = eq + 1;
        while (*v == ' ' || *v == '\t') ++v;
        size_t n = strlen(v);
        while (n && (v[n - 1] == '\n' || v[n - 1] == '\r' || v[n - 1] == ' ' || v[n - 1] == '\t')) v[--n] = '\0';
        if (!config_set(S, p, v)) fprintf(stderr, "terra_headless: config: bad value '%s' for %s (default kept)\n", v, p);
    }
    fclose(f);
    return 1;
}

static void on_tile(const char* what, size_t x, size_t y, size_t w, size_t h, size_t done, size_t total) {
    printf("terra_headless: tile %s x %zu y %zu %zux%zu (%zu/%zu)\n", what, x, y, w, h, done, total);
}

int main(int argc, char** argv) {
    const char* out = "terra.pfm";
    size_t passes = 1;
    int dry = 0, progress = 0;
    const char* builtin = "cornell";
    int devices[16], n_devices = 0;
    Settings S;
    memset(&S, 0, sizeof S);
    S.width = 256; S.height = 256; S.spp = 16; S.bounces = 8;
    S.integrator = kTerraIntegratorSimple; S.tonemap = kTerraTonemappingOperatorNone; S.sampler = kTerraSamplingMethodRandom;
    S.jitter = 0.5f; S.fov = 45.f; S.exposure = 1.f; S.gamma_ = 2.2f;
    { const float c[6] = { 0.5f, 0.5f, -1.4f, 0, 0, 1 }; memcpy(S.cam, c, sizeof c); S.up[1] = 1.f; }
    for (int i = 1; i < argc; ++i) {
        const char* a = argv[i];
#define NEED(n) do { if (i + (n) >= argc) { fprintf(stderr, "terra_headless: %s needs %d value(s)\n", a, (n)); return 2; } } while (0)
        if (!strcmp(a, "--obj")) { NEED(1); snprintf(S.obj, sizeof S.obj, "%s", argv[++i]); S.have_obj = 1; }
        else if (!strcmp(a, "--config")) { NEED(1); if (!load_config(&S, argv[++i])) return 1; }
        else if (!strcmp(a, "--out")) { NEED(1); out = argv[++i]; }
        else if (!strcmp(a, "--width")) { NEED(1); S.width = strtoul(argv[++i], NULL, 10); }
        else if (!strcmp(a, "--height")) { NEED(1); S.height = strtoul(argv[++i], NULL, 10); }
        else if (!strcmp(a, "--spp")) { NEED(1); S.spp = strtoul(argv[++i], NULL, 10); }
        else if (!strcmp(a, "--bounces")) { NEED(1); S.bounces = strtoul(argv[++i], NULL, 10); }
        else if (!strcmp(a, "--tile")) { NEED(1); S.tile = strtoul(argv[++i], NULL, 10); }
        else if (!strcmp(a, "--passes")) { NEED(1); passes = strtoul(argv[++i], NULL, 10); }
        else if (!strcmp(a, "--integrator")) { NEED(1); S.integrator = pick(argv[++i], integ, 7, "integrator"); }
        else if (!strcmp(a, "--tonemap")) { NEED(1); S.tonemap = pick(argv[++i], tones, 5, "tonemap operator"); }
        else if (!strcmp(a, "--jitter")) { NEED(1); S.jitter = strtof(argv[++i], NULL); }
        else if (!strcmp(a, "--fov")) { NEED(1); S.fov = strtof(argv[++i], NULL); }
        else if (!strcmp(a, "--exposure")) { NEED(1); S.exposure = strtof(argv[++i], NULL); }
        else if (!strcmp(a, "--gamma")) { NEED(1); S.gamma_ = strtof(argv[++i], NULL); }
        else if (!strcmp(a, "--env")) { NEED(3); for (int k = 0; k < 3; ++k) S.env[k] = strtof(argv[++i], NULL); }
        else if (!strcmp(a, "--camera")) { NEED(6); for (int k = 0; k < 6; ++k) S.cam[k] = strtof(argv[++i], NULL); S.have_cam = 1; }
        else if (!strcmp(a, "--scene")) { NEED(1); builtin = argv[++i]; }
#ifdef TERRA_HEADLESS_B200
        else if (!strcmp(a, "--devices")) {      /* "0,1,2,3": the scene is replicated on these GPUs and full-frame calls are split over them */
            NEED(1);
            for (const char* q = argv[++i]; *q && n_devices < 16;) { char* end; long v = strtol(q, &end, 10); if (end == q) break; devices[n_devices++] = (int)v; q = *end == ',' ? end + 1 : end; }
        }
#endif
        else if (!strcmp(a, "--progress")) progress = 1;
        else if (!strcmp(a, "--dry-run")) dry = 1;
        else { fprintf(stderr, "usage: terra_headless [--config satellite.config] [--obj f.obj] [--width W] [--height H] [--spp N] [--bounces N]\n"
                               "       [--integrator simple|direct|mis|mono|depth|normals|misweights] [--jitter J] [--camera px py pz dx dy dz] [--fov deg]\n"
                               "       [--tonemap none|linear|reinhard|filmic|uncharted2] [--exposure E] [--gamma G] [--env r g b] [--tile N] [--passes N]\n"
                               "       [--progress] [--out image.pfm|.ppm|.png|.hdr] [--dry-run]   (later options override earlier ones)\n"); return strcmp(a, "--help") ? 2 : 0; }
    }
    const char* obj = S.have_obj ? S.obj : NULL;
    const size_t width = S.width, height = S.height, spp = S.spp, bounces = S.bounces, tile = S.tile;
    const int integrator = S.integrator, tonemap = S.tonemap;
    const float jitter = S.jitter, fov = S.fov, exposure = S.exposure, gamma_ = S.gamma_;
    float* env = S.env; float* cam = S.cam;
    GroupVec groups = {0}; MtlVec mtls = {0};
    if (obj) { if (!load_obj(obj, &groups, &mtls)) return 1; }
    else if (!strcmp(builtin, "field")) { builtin_field(&groups, &mtls, cam); S.have_cam = 1; }
    else builtin_cornell(&groups, &mtls);
    if (groups.n == 0) { fprintf(stderr, "terra_headless: no triangles\n"); return 1; }
    if (groups.n > 64) { fprintf(stderr, "terra_headless: %zu material groups; Terra supports 64 objects (SURVEY Q14)\n", groups.n); return 1; }

    /* ---- scene: one TerraObject per material group (satellite/src/Scene.cpp:147-233) */
    HTerraScene scene = terra_scene_create();
#ifdef TERRA_HEADLESS_B200
    if (n_devices) terra_b200_scene_set_devices(scene, devices, n_devices);
#else
    (void)devices; (void)n_devices;
#endif
    size_t total = 0;
    float lo[3] = { 1e30f, 1e30f, 1e30f }, hi[3] = { -1e30f, -1e30f, -1e30f };
    for (size_t g = 0; g < groups.n; ++g) {
        const Group* G = &groups.v[g];
        const size_t n = G->pos.n / 9;
        if (n == 0) continue;
        TerraObject* o = terra_scene_add_object(scene, n);
        memcpy(o->triangles, G->pos.v, n * 9 * sizeof(float));
        for (size_t i = 0; i < n; ++i) {
            memcpy(&o->properties[i].normal_a, &G->nrm.v[i * 9], 9 * sizeof(float));
            memcpy(&o->properties[i].texcoord_a, &G->tex.v[i * 6], 6 * sizeof(float));
        }
        for (size_t i = 0; i < n * 3; ++i) for (int k = 0; k < 3; ++k) { const float v = G->pos.v[i * 3 + k]; if (v < lo[k]) lo[k] = v; if (v > hi[k]) hi[k] = v; }
        Mtl def; memset(&def, 0, sizeof def); def.kd[0] = def.kd[1] = def.kd[2] = 0.7f; def.ns = 10.f;
        const Mtl* m = (G->mtl >= 0 && (size_t)G->mtl < mtls.n) ? &mtls.v[G->mtl] : &def;
        TerraFloat3 v3;
        o->material.ior = 1.5f;
        o->material.enable_bump_map_attr = 0;
        o->material.enable_normal_map_attr = 0;
        v3.x = m->ke[0]; v3.y = m->ke[1]; v3.z = m->ke[2];
        terra_attribute_init_constant(&o->material.emissive, &v3);
        const int specular = (m->ks[0] + m->ks[1] + m->ks[2]) > 0.f && m->illum >= 2;
        if (specular) {                                       /* APOLLO_SPECULAR -> Phong, Scene.cpp:193-213 */
            v3.x = m->kd[0]; v3.y = m->kd[1]; v3.z = m->kd[2];
            terra_attribute_init_constant(&o->material.attributes[TERRA_PHONG_ALBEDO], &v3);
            v3.x = m->ks[0]; v3.y = m->ks[1]; v3.z = m->ks[2];
            terra_attribute_init_constant(&o->material.attributes[TERRA_PHONG_SPECULAR_COLOR], &v3);
            v3.x = v3.y = v3.z = m->ns;
            terra_attribute_init_constant(&o->material.attributes[TERRA_PHONG_SPECULAR_INTENSITY], &v3);
            v3.x = v3.y = v3.z = 0.f;
            terra_attribute_init_constant(&o->material.attributes[TERRA_PHONG_SAMPLE_PICK], &v3);
            o->material.attributes_count = TERRA_PHONG_END;
            terra_bsdf_phong_init(&o->material.bsdf);
        } else {                                              /* everything else -> Diffuse, Scene.cpp:216-230 */
            v3.x = m->kd[0]; v3.y = m->kd[1]; v3.z = m->kd[2];
            terra_attribute_init_constant(&o->material.attributes[TERRA_DIFFUSE_ALBEDO], &v3);
            o->material.attributes_count = TERRA_DIFFUSE_END;
            terra_bsdf_diffuse_init(&o->material.bsdf);
        }
        total += n;
    }

    /* ---- options (written through terra_scene_get_options, applied at commit: Terra.c:182) */
    TerraSceneOptions* opt = terra_scene_get_options(scene);
    { TerraFloat3 e = { env[0], env[1], env[2] }; terra_attribute_init_constant(&opt->environment_map, &e); }
    opt->tonemapping_operator = (TerraTonemappingOperator)tonemap;
    opt->accelerator = kTerraAcceleratorBVH;
    opt->sampling_method = (TerraSamplingMethod)S.sampler;
    opt->integrator = (TerraIntegrator)integrator;
    opt->subpixel_jitter = jitter;
    opt->samples_per_pixel = spp;
    opt->bounces = bounces;
    opt->strata = 4;
    opt->manual_exposure = exposure;
    opt->gamma = gamma_;

    TerraCamera camera;
    if (!S.have_cam && obj) {                                   /* frame the model from the front */
        const float cx = 0.5f * (lo[0] + hi[0]), cy = 0.5f * (lo[1] + hi[1]), ext = fmaxf(hi[0] - lo[0], fmaxf(hi[1] - lo[1], hi[2] - lo[2]));
        cam[0] = cx; cam[1] = cy; cam[2] = lo[2] - 1.5f * ext; cam[3] = 0; cam[4] = 0; cam[5] = 1;
    }
    camera.position.x = cam[0]; camera.position.y = cam[1]; camera.position.z = cam[2];
    camera.direction.x = cam[3]; camera.direction.y = cam[4]; camera.direction.z = cam[5];
    camera.up.x = S.up[0]; camera.up.y = S.up[1]; camera.up.z = S.up[2];
    camera.fov = fov;

    printf("terra_headless: %zu triangles in %zu objects, %zu materials; %zux%zu, %zu spp x %zu passes, %zu bounces, integrator %s, %s\n",
           total, terra_scene_count_objects(scene), mtls.n, width, height, spp, passes, bounces, integ[integrator],
           tile ? "tiled calls" : "one full-frame call");
    if (dry) { terra_scene_destroy(scene); return 0; }

    struct timespec t0, t1, t2;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    terra_scene_commit(scene);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    TerraFramebuffer fb;
    if (!terra_framebuffer_create(&fb, width, height)) { fprintf(stderr, "terra_headless: framebuffer %zux%zu\n", width, height); return 1; }
    double last_pass_s = 0;
    for (size_t p = 0; p < passes; ++p) {                     /* progressive: results[] accumulate across calls, Terra.c:570-572 */
        struct timespec p0, p1;
        clock_gettime(CLOCK_MONOTONIC, &p0);
        if (tile == 0) terra_render(&camera, scene, &fb, 0, 0, width, height);
        else {                                                /* the tile grid of Renderer.cpp:316-349, begin / end events as :70-98 */
            const size_t n_tiles = ((width + tile - 1) / tile) * ((height + tile - 1) / tile);
            size_t done = 0;
            for (size_t y = 0; y < height; y += tile)
                for (size_t x = 0; x < width; x += tile) {
                    const size_t tw = x + tile > width ? width - x : tile, th = y + tile > height ? height - y : tile;
                    if (progress) on_tile("begin", x, y, tw, th, done, n_tiles);
                    terra_render(&camera, scene, &fb, x, y, tw, th);
                    ++done;
                    if (progress) on_tile("end", x, y, tw, th, done, n_tiles);
                }
        }
        clock_gettime(CLOCK_MONOTONIC, &p1);
        last_pass_s = (p1.tv_sec - p0.tv_sec) + 1e-9 * (p1.tv_nsec - p0.tv_nsec);
        if (passes > 1) printf("terra_headless: pass %zu: %.4f s (%.2f Mpaths/s)\n", p + 1, last_pass_s, (double)width * height * spp / last_pass_s / 1e6);
    }
    clock_gettime(CLOCK_MONOTONIC, &t2);
    const double commit_s = (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec);
    const double render_s = (t2.tv_sec - t1.tv_sec) + 1e-9 * (t2.tv_nsec - t1.tv_nsec);
    double mean[3] = { 0, 0, 0 };
    for (size_t i = 0; i < width * height; ++i) {
        const TerraRawIntegrationResult* r = &fb.results[i];
        const double inv = r->samples > 0 ? 1.0 / r->samples : 0.0;
        mean[0] += r->acc.x * inv; mean[1] += r->acc.y * inv; mean[2] += r->acc.z * inv;
    }
    printf("terra_headless: commit %.3f s, render %.3f s (%.2f Mpaths/s), mean radiance %.6f %.6f %.6f\n", commit_s, render_s,
           (double)width * height * spp * passes / render_s / 1e6, mean[0] / (width * height), mean[1] / (width * height), mean[2] / (width * height));
    const int ok = save_image(out, &fb, tonemap == kTerraTonemappingOperatorNone);
    if (ok) printf("terra_headless: wrote %s\n", out);
    terra_framebuffer_destroy(&fb);
    terra_scene_destroy(scene);
    return ok ? 0 : 1;
}
