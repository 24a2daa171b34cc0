// Host-side schedule of HybridSuperResolutionModel.forward (reference src/architecture.py:74-82; RRDBNet / HAT per
// SURVEY.md App. A): parameter table in state-dict naming, weight packing, workspace carving and the launch sequence.
// Stream-ordered, no allocation, no synchronisation: the whole forward is CUDA-graph capturable.
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.h"
#include "ops.h"

namespace srb {

enum PKind { PK_F32 = 0, PK_CONVW = 1, PK_I64 = 2 };

struct Param {
    std::string name;
    int64_t numel;
    int kind;
    int cout, cin, taps;  // PK_CONVW only
    size_t off, bytes;    // slot in the packed buffer
};

struct ModelDesc {
    std::vector<Param> params;
    std::unordered_map<std::string, int> index;
    size_t bias_sa_off[24];
    size_t bias_oca_off[4];
    size_t mlp_w1f_off[28], mlp_b1f_off[28];  // norm2-folded fc1 (24 HAB then 4 OCAB), see fold_ln_linear
    size_t packed_bytes;
};

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

static void build_desc(ModelDesc& m) {
    size_t off = 0;
    auto add = [&](const std::string& name, int64_t numel, int kind, int cout = 0, int cin = 0, int taps = 0) {
        Param p;
        p.name = name; p.numel = numel; p.kind = kind; p.cout = cout; p.cin = cin; p.taps = taps;
        p.bytes = kind == PK_F32 ? numel * 4 : kind == PK_I64 ? numel * 8 : numel * 2;
        p.off = off;
        off = align256(off + p.bytes);
        m.index[name] = (int)m.params.size();
        m.params.push_back(p);
    };
    auto conv_tc = [&](const std::string& base, int cout, int cin, int taps) {
        add(base + ".weight", (int64_t)cout * cin * taps, PK_CONVW, cout, cin, taps);
        add(base + ".bias", cout, PK_F32);
    };
    auto conv_f32 = [&](const std::string& base, int cout, int cin, int taps) {
        add(base + ".weight", (int64_t)cout * cin * taps, PK_F32);
        add(base + ".bias", cout, PK_F32);
    };
    auto ln = [&](const std::string& base) {
        add(base + ".weight", 96, PK_F32);
        add(base + ".bias", 96, PK_F32);
    };
    char buf[160];
    // ---- stage1: RRDBNet(1,1,scale=2,nf=64,nb=23,gc=32)  (App. A.1) ----
    conv_f32("stage1.conv_first", 64, 4, 9);
    for (int i = 0; i < 23; ++i)
        for (int j = 1; j <= 3; ++j)
            for (int k = 1; k <= 5; ++k) {
                snprintf(buf, sizeof(buf), "stage1.body.%d.rdb%d.conv%d", i, j, k);
                conv_tc(buf, k == 5 ? 64 : 32, 64 + 32 * (k - 1), 9);
            }
    conv_tc("stage1.conv_body", 64, 64, 9);
    conv_tc("stage1.conv_up1", 64, 64, 9);
    conv_tc("stage1.conv_up2", 64, 64, 9);
    conv_tc("stage1.conv_hr", 64, 64, 9);
    conv_f32("stage1.conv_last", 1, 64, 9);
    // ---- stage2: HAT (App. A.2 / A.3) ----
    add("stage2.relative_position_index_SA", 64 * 64, PK_I64);
    add("stage2.relative_position_index_OCA", 64 * 144, PK_I64);
    conv_f32("stage2.conv_first", 96, 1, 9);
    ln("stage2.patch_embed.norm");
    for (int l = 0; l < 4; ++l) {
        snprintf(buf, sizeof(buf), "stage2.layers.%d.residual_group.overlap_attn.", l);
        const std::string oca = buf;
        for (int j = 0; j < 6; ++j) {
            snprintf(buf, sizeof(buf), "stage2.layers.%d.residual_group.blocks.%d.", l, j);
            const std::string b = buf;
            ln(b + "norm1");
            add(b + "attn.relative_position_bias_table", 225 * 6, PK_F32);
            conv_tc(b + "attn.qkv", 288, 96, 1);
            conv_tc(b + "attn.proj", 96, 96, 1);
            conv_tc(b + "conv_block.cab.0", 32, 96, 9);
            conv_tc(b + "conv_block.cab.2", 96, 32, 9);
            conv_f32(b + "conv_block.cab.3.attention.1", 3, 96, 1);
            conv_f32(b + "conv_block.cab.3.attention.3", 96, 3, 1);
            ln(b + "norm2");
            conv_tc(b + "mlp.fc1", 192, 96, 1);
            conv_tc(b + "mlp.fc2", 96, 192, 1);
        }
        snprintf(buf, sizeof(buf), "stage2.layers.%d.residual_group.overlap_attn.", l);
        const std::string b = buf;
        add(b + "relative_position_bias_table", 361 * 6, PK_F32);
        ln(b + "norm1");
        conv_tc(b + "qkv", 288, 96, 1);
        conv_tc(b + "proj", 96, 96, 1);
        ln(b + "norm2");
        conv_tc(b + "mlp.fc1", 192, 96, 1);
        conv_tc(b + "mlp.fc2", 96, 192, 1);
        snprintf(buf, sizeof(buf), "stage2.layers.%d.conv", l);
        conv_tc(buf, 96, 96, 9);
    }
    ln("stage2.norm");
    conv_tc("stage2.conv_after_body", 96, 96, 9);
    conv_tc("stage2.conv_before_upsample.0", 64, 96, 9);
    conv_tc("stage2.upsample.0", 256, 64, 9);
    conv_f32("stage2.conv_last", 1, 64, 9);
    // ---- derived: gathered relative-position bias tables ----
    for (int i = 0; i < 24; ++i) { m.bias_sa_off[i] = off; off = align256(off + 6 * 64 * 64 * 2); }  // bf16, fragment order
    for (int i = 0; i < 4; ++i) { m.bias_oca_off[i] = off; off = align256(off + 6 * 64 * 144 * 4); }
    for (int i = 0; i < 28; ++i) {
        m.mlp_w1f_off[i] = off; off = align256(off + 192 * 96 * 2);
        m.mlp_b1f_off[i] = off; off = align256(off + 192 * 4);
    }
    m.packed_bytes = off;
}

static const ModelDesc& desc() {
    static ModelDesc m;
    static std::once_flag once;
    std::call_once(once, [] { build_desc(m); });
    return m;
}

int model_param_count() { return (int)desc().params.size(); }
const char* model_param_name(int i) {
    const ModelDesc& m = desc();
    return (i >= 0 && i < (int)m.params.size()) ? m.params[i].name.c_str() : nullptr;
}
int64_t model_param_numel(int i) {
    const ModelDesc& m = desc();
    return (i >= 0 && i < (int)m.params.size()) ? m.params[i].numel : -1;
}
size_t model_packed_bytes() { return desc().packed_bytes; }

int model_pack(const void* const* params, int n, void* packed, size_t packed_bytes, cudaStream_t s) {
    const ModelDesc& m = desc();
    if (n != (int)m.params.size()) { set_error("pack_weights: expected %d tensors, got %d", (int)m.params.size(), n); return SRB_E_ARG; }
    if (packed_bytes < m.packed_bytes) { set_error("pack_weights: packed buffer too small (%zu < %zu)", packed_bytes, m.packed_bytes); return SRB_E_WORKSPACE; }
    uint8_t* base = (uint8_t*)packed;
    for (int i = 0; i < n; ++i) {
        const Param& p = m.params[i];
        if (!params[i]) { set_error("pack_weights: null pointer for %s", p.name.c_str()); return SRB_E_ARG; }
        if (p.kind == PK_CONVW) {
            SRB_TRY(pack_conv_weight((const float*)params[i], (bf16*)(base + p.off), p.cout, p.cin, p.taps, s));
        } else {
            SRB_CUDA(cudaMemcpyAsync(base + p.off, params[i], p.bytes, cudaMemcpyDeviceToDevice, s));
        }
    }
    const int64_t* idx_sa = (const int64_t*)(base + m.params[m.index.at("stage2.relative_position_index_SA")].off);
    const int64_t* idx_oca = (const int64_t*)(base + m.params[m.index.at("stage2.relative_position_index_OCA")].off);
    char buf[160];
    for (int l = 0; l < 4; ++l) {
        for (int j = 0; j < 6; ++j) {
            snprintf(buf, sizeof(buf), "stage2.layers.%d.residual_group.blocks.%d.attn.relative_position_bias_table", l, j);
            const float* tab = (const float*)(base + m.params[m.index.at(buf)].off);
            SRB_TRY(expand_rel_bias_bf16(tab, 225, idx_sa, 64, (bf16*)(base + m.bias_sa_off[l * 6 + j]), 1.4426950408889634f, s));
        }
        snprintf(buf, sizeof(buf), "stage2.layers.%d.residual_group.overlap_attn.relative_position_bias_table", l);
        const float* tab = (const float*)(base + m.params[m.index.at(buf)].off);
        SRB_TRY(expand_rel_bias_bf16(tab, 361, idx_oca, 144, (bf16*)(base + m.bias_oca_off[l]), 1.4426950408889634f, s));
    }
    auto src = [&](const std::string& name) { return (const float*)params[m.index.at(name)]; };
    // OCAB: the q rows of qkv carry head_dim^-0.5 * log2(e) (the attention core works in the log2 domain)
    for (int l = 0; l < 4; ++l) {
        snprintf(buf, sizeof(buf), "stage2.layers.%d.residual_group.overlap_attn.qkv", l);
        const std::string b = buf;
        SRB_TRY(pack_linear_scaled(src(b + ".weight"), src(b + ".bias"), (bf16*)(base + m.params[m.index.at(b + ".weight")].off),
                                   (float*)(base + m.params[m.index.at(b + ".bias")].off), 288, 96, 96,
                                   0.25f * 1.4426950408889634f, s));
    }
    // norm2 folded into mlp.fc1 (from the fp32 originals: one bf16 rounding)
    for (int i = 0; i < 28; ++i) {
        if (i < 24) snprintf(buf, sizeof(buf), "stage2.layers.%d.residual_group.blocks.%d.", i / 6, i % 6);
        else snprintf(buf, sizeof(buf), "stage2.layers.%d.residual_group.overlap_attn.", i - 24);
        const std::string b = buf;
        SRB_TRY(fold_ln_linear(src(b + "mlp.fc1.weight"), src(b + "mlp.fc1.bias"), src(b + "norm2.weight"),
                               src(b + "norm2.bias"), (bf16*)(base + m.mlp_w1f_off[i]),
                               (float*)(base + m.mlp_b1f_off[i]), 192, 96, s));
    }
    return SRB_OK;
}

// ------------------------------------------------------------------------------------------------------------------
// workspace (HBM layout): all activations NHWC ("pixel-major"), bf16 unless noted
// ------------------------------------------------------------------------------------------------------------------
struct Workspace {
    // stage 1 (trunk at h x w = H/2 x W/2)
    bf16* cb[2];   // [B,h,w,192] dense-block buffers (ping-pong): [x | x1 | x2 | x3 | x4]
    float* t[2];   // [B,h,w,64] fp32 residual trunk (running / RRDB input)
    float* f0;     // [B,h,w,64] fp32 conv_first output (long skip)
    bf16* fe;      // [B,h,w,64] feat + conv_body(body(feat))
    bf16 *u1, *v1; // [B,H,W,64] nearest x2 / conv_up1
    bf16 *u2, *v2, *v3;  // [B,2H,2W,64] nearest x2 / conv_up2 / conv_hr
    float *s1a, *s1b;    // [B,2H,2W] fp32 stage-1 image before / after s1
    // stage 2 (tokens at R x Rw = 2H x 2W)
    bf16 *f0h, *g, *y, *xn, *ao, *c2;  // [ntok,96]
    bf16* c1;       // [ntok,32]
    bf16* qkv;      // [ntok,288]
    bf16* hd;       // [ntok,192]
    float *pool, *cs;  // [B,96]
    bf16* d1;       // [ntok,64]
    bf16* d2;       // [4*ntok,64]
    float *o1, *o2; // [B,4H,4W] fp32
    unsigned int* chain_ctr;  // [23][16] grid-barrier counters of the RRDB chain kernels (zeroed per forward)
};

static size_t carve(Workspace& w, uint8_t* base, int B, int H, int W) {
    size_t off = 0;
    auto take = [&](size_t bytes) {
        uint8_t* p = base ? base + off : nullptr;
        off = align256(off + bytes);
        return p;
    };
    const size_t h = H / 2, ww = W / 2;
    const size_t trunk = (size_t)B * h * ww;
    for (int i = 0; i < 2; ++i) w.cb[i] = (bf16*)take(trunk * 192 * 2);
    for (int i = 0; i < 2; ++i) w.t[i] = (float*)take(trunk * 64 * 4);
    w.f0 = (float*)take(trunk * 64 * 4);
    w.fe = (bf16*)take(trunk * 64 * 2);
    const size_t p1 = (size_t)B * H * W, p2 = p1 * 4, p4 = p1 * 16;
    w.u1 = (bf16*)take(p1 * 64 * 2);
    w.v1 = (bf16*)take(p1 * 64 * 2);
    w.u2 = (bf16*)take(p2 * 64 * 2);
    w.v2 = (bf16*)take(p2 * 64 * 2);
    w.v3 = (bf16*)take(p2 * 64 * 2);
    w.s1a = (float*)take(p2 * 4);
    w.s1b = (float*)take(p2 * 4);
    const size_t ntok = p2;
    w.f0h = (bf16*)take(ntok * 96 * 2);
    w.g = (bf16*)take(ntok * 96 * 2);
    w.y = (bf16*)take(ntok * 96 * 2);
    w.xn = (bf16*)take(ntok * 96 * 2);
    w.ao = (bf16*)take(ntok * 96 * 2);
    w.c2 = (bf16*)take(ntok * 96 * 2);
    w.c1 = (bf16*)take(ntok * 32 * 2);
    w.qkv = (bf16*)take(ntok * 288 * 2);
    w.hd = (bf16*)take(ntok * 192 * 2);
    w.pool = (float*)take((size_t)B * cab_pool_slabs(2 * H, 2 * W) * 96 * 4);  // per-tile partial sums
    w.cs = (float*)take((size_t)B * 96 * 4);
    w.d1 = (bf16*)take(ntok * 64 * 2);
    w.d2 = (bf16*)take(p4 * 64 * 2);
    w.o1 = (float*)take(p4 * 4);
    w.o2 = (float*)take(p4 * 4);
    w.chain_ctr = (unsigned int*)take(23 * 16 * sizeof(unsigned int));
    return off;
}

static int check_geometry(int B, int H, int W, int stage_mask) {
    if (B <= 0 || H <= 0 || W <= 0) { set_error("forward: bad geometry B=%d H=%d W=%d", B, H, W); return SRB_E_SHAPE; }
    // The reference accepts what pixel_unshuffle(2) and the 8 x 8 HAT windows accept: stage 1 runs its trunk at H/2 x W/2,
    // stage 2 its tokens at 2H x 2W, which must tile into windows -> H, W multiples of 4 (src/architecture.py:74-82).
    // Tile-ragged sizes are handled inside the kernels (TMA zero fill / clipped stores, masked epilogues); the RRDB
    // chain kernel needs a 16 x 32-tiled trunk and otherwise gives way to one launch per conv.
    if ((stage_mask & 1) && (H % 4 || W % 4)) {
        set_error("forward: input H=%d W=%d must be multiples of 4", H, W);
        return SRB_E_SHAPE;
    }
    if (stage_mask == 2 && (H % 8 || W % 8)) {
        set_error("forward(stage 2 only): input H=%d W=%d must be multiples of the HAT window 8", H, W);
        return SRB_E_SHAPE;
    }
    return SRB_OK;
}

size_t model_workspace_bytes(int B, int H, int W) {
    Workspace w;
    return carve(w, nullptr, B, H, W);
}

// ------------------------------------------------------------------------------------------------------------------
// forward schedule
// ------------------------------------------------------------------------------------------------------------------
enum ProfClass { PC_CONV3 = 0, PC_LINEAR = 1, PC_ATTN = 2, PC_LN = 3, PC_OTHER = 4, PC_MLP = 5, PC_OCAB = 6, PC_CHAIN = 7, PC_COUNT = 8 };
static_assert(PC_COUNT == SRB_PROF_CLASSES, "profile classes");
struct ProfRec {
    int cls;
    double flops;
    cudaEvent_t e0, e1;
};

struct Ctx {
    std::vector<ProfRec>* prof = nullptr;  // per-launch CUDA events on the launching stream (profile mode only)
    const ModelDesc* m;
    const uint8_t* packed;
    cudaStream_t s;
    bool dry;     // count launches only
    int launches;
    const void* P(const std::string& name) const { return packed + m->params[m->index.at(name)].off; }
    const float* F(const std::string& name) const { return (const float*)P(name); }
    const bf16* Wt(const std::string& name) const { return (const bf16*)P(name); }
};

static void prof_begin(Ctx& c, int cls, double flops) {
    if (!c.prof || c.dry) return;
    ProfRec r;
    r.cls = cls; r.flops = flops;
    cudaEventCreate(&r.e0);
    cudaEventCreate(&r.e1);
    cudaEventRecord(r.e0, c.s);
    c.prof->push_back(r);
}
static void prof_end(Ctx& c) {
    if (!c.prof || c.dry) return;
    cudaEventRecord(c.prof->back().e1, c.s);
}

// one tensor-core conv / linear layer named by its state-dict prefix
static ConvArgs make_layer(const Ctx& c, const std::string& name, const bf16* in, int in_pitch, int B, int H, int W, int act,
                           float slope, bf16* out, int out_pitch, float* out32 = nullptr, float alpha1 = 1.f,
                           const void* res1 = nullptr, int res1_f32 = 0, int res1_pitch = 0, float alpha2 = 1.f,
                           const void* res2 = nullptr, int res2_f32 = 0, int res2_pitch = 0,
                           const float* res2_cs = nullptr, int pixel_shuffle = 0) {
    const Param& p = c.m->params[c.m->index.at(name + ".weight")];
    ConvArgs a;
    memset(&a, 0, sizeof(a));
    a.B = B; a.H = H; a.W = W; a.Cin = p.cin; a.Cout = p.cout; a.taps = p.taps;
    a.in = in; a.in_pitch = in_pitch;
    a.w = (const bf16*)(c.packed + p.off);
    a.bias = c.F(name + ".bias");
    a.act = act; a.slope = slope;
    a.alpha1 = alpha1; a.res1 = res1; a.res1_f32 = res1_f32; a.res1_pitch = res1_pitch;
    a.alpha2 = alpha2; a.res2 = res2; a.res2_f32 = res2_f32; a.res2_pitch = res2_pitch; a.res2_cs = res2_cs;
    a.out = out; a.out_pitch = out_pitch; a.out32 = out32; a.out32_pitch = 64;
    a.pixel_shuffle = pixel_shuffle;
    return a;
}
static double layer_flops(const ConvArgs& a) { return 2.0 * a.B * a.H * a.W * (double)a.Cout * a.Cin * a.taps; }

static int tc_layer(Ctx& c, const std::string& name, const bf16* in, int in_pitch, int B, int H, int W, int act,
                    float slope, bf16* out, int out_pitch, float* out32 = nullptr, float alpha1 = 1.f,
                    const void* res1 = nullptr, int res1_f32 = 0, int res1_pitch = 0, float alpha2 = 1.f,
                    const void* res2 = nullptr, int res2_f32 = 0, int res2_pitch = 0, const float* res2_cs = nullptr,
                    int pixel_shuffle = 0, float* colsum = nullptr) {
    ++c.launches;
    if (c.dry) return SRB_OK;
    ConvArgs a = make_layer(c, name, in, in_pitch, B, H, W, act, slope, out, out_pitch, out32, alpha1, res1,
                                  res1_f32, res1_pitch, alpha2, res2, res2_f32, res2_pitch, res2_cs, pixel_shuffle);
    a.colsum = colsum;
    prof_begin(c, a.taps == 9 ? PC_CONV3 : PC_LINEAR, layer_flops(a));
    const int rc = conv_tc_launch(a, c.s);
    prof_end(c);
    return rc;
}

#define LAUNCH(c, cls, flops, expr)     \
    do {                                \
        ++(c).launches;                 \
        if (!(c).dry) {                 \
            prof_begin(c, cls, flops);  \
            const int rc__ = (expr);    \
            prof_end(c);                \
            if (rc__ != SRB_OK) return rc__; \
        }                               \
    } while (0)

// RRDBNet.forward (App. A.1): x [B,1,H,W] -> s1a fp32 [B,2H,2W]
static int run_stage1(Ctx& c, Workspace& w, const void* x, int x_dtype, int B, int H, int W) {
    const int h = H / 2, ww = W / 2;
    LAUNCH(c, PC_OTHER, 0.0, rrdb_conv_first(x, x_dtype, c.F("stage1.conv_first.weight"), c.F("stage1.conv_first.bias"), w.cb[0], 192,
                              w.t[0], w.f0, B, h, ww, c.s));
    int cur = 0;
    char buf[96];
    // One RRDB (15 dependent convs) = one persistent chain kernel when the trunk tiles as 16 x 32 (conv_chain.cu);
    // narrower trunks fall back to one launch per conv.
    const bool chain = (h % 16 == 0) && (ww % 32 == 0) && conv_chain_max_layers() >= 15;
    if (chain && !c.dry) SRB_CUDA(cudaMemsetAsync(w.chain_ctr, 0, 23 * 16 * sizeof(unsigned int), c.s));
    for (int i = 0; i < 23; ++i) {
        ConvArgs la[15];
        int nl = 0;
        for (int j = 1; j <= 3; ++j) {
            bf16* cb = w.cb[cur];
            bf16* nxt = w.cb[cur ^ 1];
            for (int k = 1; k <= 5; ++k) {
                snprintf(buf, sizeof(buf), "stage1.body.%d.rdb%d.conv%d", i, j, k);
                if (c.dry) { ++nl; continue; }
                if (k < 5)
                    la[nl++] = make_layer(c, buf, cb, 192, B, h, ww, ACT_LRELU, 0.2f, cb + 64 + 32 * (k - 1), 192);
                // fp32 trunk: t[0] holds the RRDB input, t[1] the running RDB output
                else if (j == 1)
                    la[nl++] = make_layer(c, buf, cb, 192, B, h, ww, ACT_NONE, 0.f, nxt, 192, w.t[1], 0.2f, w.t[0], 1, 64);
                else if (j == 2)
                    la[nl++] = make_layer(c, buf, cb, 192, B, h, ww, ACT_NONE, 0.f, nxt, 192, w.t[1], 0.2f, w.t[1], 1, 64);
                else  // (x5*0.2 + x_rdb3)*0.2 + x_rrdb
                    la[nl++] = make_layer(c, buf, cb, 192, B, h, ww, ACT_NONE, 0.f, nxt, 192, w.t[0], 0.2f, w.t[1], 1, 64,
                                          0.2f, w.t[0], 1, 64);
            }
            cur ^= 1;
        }
        if (chain) {
            ++c.launches;
            if (!c.dry) {
                double fl = 0;
                for (int q = 0; q < nl; ++q) fl += layer_flops(la[q]);
                prof_begin(c, PC_CHAIN, fl);
                const int rc = conv_chain_launch(la, nl, w.chain_ctr + 16 * i, c.s);
                prof_end(c);
                if (rc != SRB_OK) return rc;
            }
        } else {
            c.launches += nl;
            if (!c.dry)
                for (int q = 0; q < nl; ++q) {
                    prof_begin(c, PC_CONV3, layer_flops(la[q]));
                    const int rc = conv_tc_launch(la[q], c.s);
                    prof_end(c);
                    if (rc != SRB_OK) return rc;
                }
        }
    }
    // feat = feat + conv_body(body(feat))
    SRB_TRY(tc_layer(c, "stage1.conv_body", w.cb[cur], 192, B, h, ww, ACT_NONE, 0.f, w.fe, 64, nullptr, 1.f, w.f0, 1, 64));
    LAUNCH(c, PC_OTHER, 0.0, upsample2x(w.fe, 64, w.u1, B, h, ww, c.s));
    SRB_TRY(tc_layer(c, "stage1.conv_up1", w.u1, 64, B, H, W, ACT_LRELU, 0.2f, w.v1, 64));
    LAUNCH(c, PC_OTHER, 0.0, upsample2x(w.v1, 64, w.u2, B, H, W, c.s));
    SRB_TRY(tc_layer(c, "stage1.conv_up2", w.u2, 64, B, 2 * H, 2 * W, ACT_LRELU, 0.2f, w.v2, 64));
    SRB_TRY(tc_layer(c, "stage1.conv_hr", w.v2, 64, B, 2 * H, 2 * W, ACT_LRELU, 0.2f, w.v3, 64));
    LAUNCH(c, PC_OTHER, 0.0, conv_cout1(w.v3, 64, c.F("stage1.conv_last.weight"), c.F("stage1.conv_last.bias"), w.s1a, B, 2 * H, 2 * W, c.s));
    return SRB_OK;
}

// HAT.forward (App. A.2): img fp32 [B,R,Rw] -> o1 fp32 [B,2R,2Rw]
static int run_stage2(Ctx& c, Workspace& w, const float* img, int B, int R, int Rw) {
    const size_t ntok = (size_t)B * R * Rw;
    const uint8_t* pk = c.packed;
    LAUNCH(c, PC_OTHER, 0.0, hat_conv_first_ln(img, c.F("stage2.conv_first.weight"), c.F("stage2.conv_first.bias"),
                                c.F("stage2.patch_embed.norm.weight"), c.F("stage2.patch_embed.norm.bias"), w.f0h, w.g,
                                B, R, Rw, c.s));
    char buf[128];
    for (int l = 0; l < 4; ++l) {
        snprintf(buf, sizeof(buf), "stage2.layers.%d.residual_group.overlap_attn.", l);
        const std::string oca = buf;
        for (int j = 0; j < 6; ++j) {
            snprintf(buf, sizeof(buf), "stage2.layers.%d.residual_group.blocks.%d.", l, j);
            const std::string b = buf;
            const bf16* xin = (j == 0) ? w.g : w.y;  // token stream entering this block (= shortcut)
            // fused attention branch: xn = LN1(x), ao = x + proj(W-MSA(xn))   (LN1, qkv, attention, proj in one kernel)
            LAUNCH(c, PC_ATTN, 2.0 * (27648 + 12288 + 9216) * (double)ntok,
                   hab_attention_tc(xin, c.F(b + "norm1.weight"), c.F(b + "norm1.bias"), c.Wt(b + "attn.qkv.weight"),
                                       c.F(b + "attn.qkv.bias"), (const bf16*)(pk + c.m->bias_sa_off[l * 6 + j]),
                                       c.Wt(b + "attn.proj.weight"), c.F(b + "attn.proj.bias"), w.xn, w.ao, B, R, Rw,
                                       (j & 1) ? 4 : 0, c.s));
            // CAB on the normalised tensor: conv 96->32, GELU, conv 32->96, channel attention
            SRB_TRY(tc_layer(c, b + "conv_block.cab.0", w.xn, 96, B, R, Rw, ACT_GELU, 0.f, w.c1, 32));
            // cab.2 also leaves per-tile channel sums in w.pool (the average pool of the channel attention, fused)
            SRB_TRY(tc_layer(c, b + "conv_block.cab.2", w.c1, 32, B, R, Rw, ACT_NONE, 0.f, w.c2, 96, nullptr, 1.f, nullptr,
                             0, 0, 1.f, nullptr, 0, 0, nullptr, 0, w.pool));
            LAUNCH(c, PC_OTHER, 0.0, cab_channel_scale(w.pool, cab_pool_slabs(R, Rw), w.cs,
                                            c.F(b + "conv_block.cab.3.attention.1.weight"),
                                            c.F(b + "conv_block.cab.3.attention.1.bias"),
                                            c.F(b + "conv_block.cab.3.attention.3.weight"),
                                            c.F(b + "conv_block.cab.3.attention.3.bias"), 0.01f, B, R * Rw, c.s));
            // y = (ao + cab * sigmoid(..) * conv_scale) + fc2(GELU(fc1(LN2(.))))  — combine, norm2 and the MLP in one kernel
            LAUNCH(c, PC_MLP, 2.0 * 36864 * (double)ntok,
                   mlp_fused(w.ao, w.c2, w.cs, (const bf16*)(pk + c.m->mlp_w1f_off[l * 6 + j]),
                             (const float*)(pk + c.m->mlp_b1f_off[l * 6 + j]), c.Wt(b + "mlp.fc2.weight"),
                             c.F(b + "mlp.fc2.bias"), w.y, ntok, (size_t)R * Rw, c.s,
                             // the last HAB of a group also emits xn = OCAB.norm1(y) from its copy-out
                             j == 5 ? c.F(oca + "norm1.weight") : nullptr, j == 5 ? c.F(oca + "norm1.bias") : nullptr,
                             j == 5 ? w.xn : nullptr));
        }
        const std::string b = oca;
        SRB_TRY(tc_layer(c, b + "qkv", w.xn, 96, B, R, Rw, ACT_NONE, 0.f, w.qkv, 288));
        LAUNCH(c, PC_OCAB, 2.0 * 2 * 64 * 144 * 16 * 6 * ((double)B * R * Rw / 64), ocab_attention(w.qkv, (const bf16*)(pk + c.m->bias_oca_off[l]), w.ao, B, R, Rw, c.s));
        SRB_TRY(tc_layer(c, b + "proj", w.ao, 96, B, R, Rw, ACT_NONE, 0.f, w.y, 96, nullptr, 1.f, w.y, 0, 96));
        LAUNCH(c, PC_MLP, 2.0 * 36864 * (double)ntok,
               mlp_fused(w.y, nullptr, nullptr, (const bf16*)(pk + c.m->mlp_w1f_off[24 + l]),
                         (const float*)(pk + c.m->mlp_b1f_off[24 + l]), c.Wt(b + "mlp.fc2.weight"),
                         c.F(b + "mlp.fc2.bias"), w.y, ntok, (size_t)R * Rw, c.s));
        // RHAG tail: g = conv(y) + g
        snprintf(buf, sizeof(buf), "stage2.layers.%d.conv", l);
        SRB_TRY(tc_layer(c, buf, w.y, 96, B, R, Rw, ACT_NONE, 0.f, w.g, 96, nullptr, 1.f, w.g, 0, 96));
    }
    LAUNCH(c, PC_LN, 0.0, layernorm96(w.g, c.F("stage2.norm.weight"), c.F("stage2.norm.bias"), w.xn, ntok, c.s));
    SRB_TRY(tc_layer(c, "stage2.conv_after_body", w.xn, 96, B, R, Rw, ACT_NONE, 0.f, w.y, 96, nullptr, 1.f, w.f0h, 0, 96));
    SRB_TRY(tc_layer(c, "stage2.conv_before_upsample.0", w.y, 96, B, R, Rw, ACT_LRELU, 0.01f, w.d1, 64));
    SRB_TRY(tc_layer(c, "stage2.upsample.0", w.d1, 64, B, R, Rw, ACT_NONE, 0.f, w.d2, 64, nullptr, 1.f, nullptr, 0, 0, 1.f,
                     nullptr, 0, 0, nullptr, 1));
    LAUNCH(c, PC_OTHER, 0.0, conv_cout1(w.d2, 64, c.F("stage2.conv_last.weight"), c.F("stage2.conv_last.bias"), w.o1, B, 2 * R, 2 * Rw, c.s));
    return SRB_OK;
}

static int run_forward(Ctx& c, const void* x, int x_dtype, void* y, int y_dtype, int B, int H, int W, void* workspace,
                       int smoothing, int stage_mask) {
    const bool stop_before_sf = stage_mask == 7;  // the caller resizes (src/architecture.py:80-81) and applies sf itself
    if (stop_before_sf) stage_mask = 3;
    Workspace w;
    memset(&w, 0, sizeof(w));
    if (stage_mask == 2) carve(w, (uint8_t*)workspace, B, H / 2, W / 2);
    else carve(w, (uint8_t*)workspace, B, H, W);
    if (stage_mask == 1) {  // RRDBNet only: y = stage-1 image [B,1,2H,2W]
        SRB_TRY(run_stage1(c, w, x, x_dtype, B, H, W));
        LAUNCH(c, PC_OTHER, 0.0, convert_out(w.s1a, y, y_dtype, (size_t)B * 4 * H * W, c.s));
        return SRB_OK;
    }
    if (stage_mask == 2) {  // HAT only: x fp32 [B,1,H,W] -> y [B,1,2H,2W]
        if (x_dtype != SRB_F32) { set_error("forward(stage 2 only) needs fp32 input"); return SRB_E_DTYPE; }
        SRB_TRY(run_stage2(c, w, (const float*)x, B, H, W));
        LAUNCH(c, PC_OTHER, 0.0, convert_out(w.o1, y, y_dtype, (size_t)B * 4 * H * W, c.s));
        return SRB_OK;
    }
    SRB_TRY(run_stage1(c, w, x, x_dtype, B, H, W));
    const float* img = w.s1a;
    if (smoothing != SRB_SMOOTH_NONE) {
        LAUNCH(c, PC_OTHER, 0.0, blur(w.s1a, w.s1b, SRB_F32, smoothing, B, 2 * H, 2 * W, c.s));
        img = w.s1b;
    }
    SRB_TRY(run_stage2(c, w, img, B, 2 * H, 2 * W));
    const size_t nout = (size_t)B * 16 * H * W;
    if (smoothing != SRB_SMOOTH_NONE && stop_before_sf) {
        LAUNCH(c, PC_OTHER, 0.0, blur(w.o1, y, y_dtype, smoothing, B, 4 * H, 4 * W, c.s));
    } else if (smoothing != SRB_SMOOTH_NONE) {
        LAUNCH(c, PC_OTHER, 0.0, blur(w.o1, w.o2, SRB_F32, smoothing, B, 4 * H, 4 * W, c.s));
        LAUNCH(c, PC_OTHER, 0.0, blur(w.o2, y, y_dtype, SRB_SMOOTH_LIGHT, B, 4 * H, 4 * W, c.s));
    } else {
        LAUNCH(c, PC_OTHER, 0.0, convert_out(w.o1, y, y_dtype, nout, c.s));
    }
    return SRB_OK;
}

int model_forward(const void* x, int x_dtype, void* y, int y_dtype, int B, int H, int W, const void* packed,
                  void* workspace, size_t workspace_bytes, int smoothing, int stage_mask, cudaStream_t s, bool dry,
                  int* launches, double* prof_ms, double* prof_flops, int* prof_counts) {
    if ((stage_mask < 1 || stage_mask > 3) && stage_mask != 7) { set_error("forward: stage_mask must be 1, 2, 3 or 7"); return SRB_E_ARG; }
    if (smoothing < 0 || smoothing > 3) { set_error("forward: bad smoothing mode %d", smoothing); return SRB_E_ARG; }
    if (y_dtype != SRB_F32 && y_dtype != SRB_U16) { set_error("forward: y dtype must be F32 or U16"); return SRB_E_DTYPE; }
    if (x_dtype != SRB_F32 && x_dtype != SRB_U16 && x_dtype != SRB_BF16) { set_error("forward: bad x dtype"); return SRB_E_DTYPE; }
    SRB_TRY(check_geometry(B, H, W, stage_mask));
    if (!dry) {
        if (!x || !y || !packed || !workspace) { set_error("forward: null pointer"); return SRB_E_ARG; }
        const size_t need = stage_mask == 2 ? model_workspace_bytes(B, H / 2, W / 2) : model_workspace_bytes(B, H, W);
        if (workspace_bytes < need) { set_error("forward: workspace too small (%zu < %zu)", workspace_bytes, need); return SRB_E_WORKSPACE; }
    }
    Ctx c;
    c.m = &desc(); c.packed = (const uint8_t*)packed; c.s = s; c.dry = dry; c.launches = 0;
    std::vector<ProfRec> recs;
    if (prof_ms && !dry) c.prof = &recs;
    int rc = run_forward(c, x, x_dtype, y, y_dtype, B, H, W, workspace, smoothing, stage_mask);
    if (launches) *launches = c.launches;
    if (c.prof) {
        // profile mode only: wait for the stream, then fold the per-launch event times into per-class totals
        for (int i = 0; i < PC_COUNT; ++i) { prof_ms[i] = 0; prof_flops[i] = 0; prof_counts[i] = 0; }
        cudaError_t e = cudaStreamSynchronize(s);
        for (ProfRec& r : recs) {
            float ms = 0.f;
            if (e == cudaSuccess && cudaEventElapsedTime(&ms, r.e0, r.e1) == cudaSuccess) {
                prof_ms[r.cls] += ms; prof_flops[r.cls] += r.flops; prof_counts[r.cls] += 1;
            }
            cudaEventDestroy(r.e0);
            cudaEventDestroy(r.e1);
        }
        if (e != cudaSuccess && rc == SRB_OK) rc = cuda_fail(e, "cudaStreamSynchronize(profile)");
    }
    return rc;
}

}  // namespace srb
