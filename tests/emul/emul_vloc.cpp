// TEST INFRASTRUCTURE (not product code): executes the phase functions of the
// V_loc CUDA kernels (q-e_b200/csrc/vloc_kernels.cuh) serially on the CPU so
// that the index / twiddle / table logic can be checked against the oracle in
// the "-m 'not gpu'" suite. Built by tests/emul/build.sh with plain g++.
#include <stdio.h>
#include <string.h>
#include <vector>
#include "vloc_setup.h"
#include "fullfft_kernels.cuh"
#include "vloc_fast.cuh"
#include "vloc_plane.cuh"
#include "vloc_plane2.cuh"

using namespace qeb;

static void upload_plan(const HostPlan& hp, FftPlan& pl) { fill_plan(hp, hp.tw.data(), hp.rev.data(), pl); }

extern "C" int emul_fft1d(int n, int sign, int nlines, const double* in, double* out, int nthr, int maxr)
{
    // nlines lines of length n, layout [line][n]; out same layout, natural order.
    HostPlan hp;
    if (!make_host_plan(n, hp, maxr)) return 1;
    FftPlan pl;
    upload_plan(hp, pl);
    std::vector<cplx> s((size_t)n * nlines);
    const cplx* ci = (const cplx*)in;
    cplx* co = (cplx*)out;
    FastDiv dnl = make_fastdiv(nlines);
    SmemTile st{s.data(), nlines};
    NoMid nm;
    if (sign > 0) {  // DIF natural in
        for (int l = 0; l < nlines; ++l)
            for (int i = 0; i < n; ++i) s[(size_t)i * nlines + l] = ci[(size_t)l * n + i];
        for (int p = 0; p < pl.npass; ++p) {
            PassDesc pd = pass_of(pl, p);
            for (int tid = 0; tid < nthr; ++tid) fft_pass<false, false>(tid, nthr, nlines, dnl, pd, pl.tw, +1.0, st, st, nm);
        }
        for (int l = 0; l < nlines; ++l)
            for (int i = 0; i < n; ++i) co[(size_t)l * n + hp.rev[i]] = s[(size_t)i * nlines + l];
    } else {  // DIT digit-reversed in
        for (int l = 0; l < nlines; ++l)
            for (int i = 0; i < n; ++i) s[(size_t)i * nlines + l] = ci[(size_t)l * n + hp.rev[i]];
        for (int p = pl.npass - 1; p >= 0; --p) {
            PassDesc pd = pass_of(pl, p);
            for (int tid = 0; tid < nthr; ++tid) fft_pass<true, false>(tid, nthr, nlines, dnl, pd, pl.tw, -1.0, st, st, nm);
        }
        for (int l = 0; l < nlines; ++l)
            for (int i = 0; i < n; ++i) co[(size_t)l * n + i] = s[(size_t)i * nlines + l];
    }
    return 0;
}

#define QEB_FOR_PLANS(X) X(SP15) X(SP25) X(SP27) X(SP64) X(SP125) X(SP135) X(SP144) X(SP200) X(SP200b) X(SP200c) X(SP200d)
typedef void (*zfn)(int, int, int, const VlocTables&, const VlocIO&, int, int, cplx*, const FastTabs&);
typedef void (*yfn)(int, int, int, const VlocTables&, int, int, int, cplx*, const FastTabs&);
typedef void (*xfn)(int, int, int, const VlocTables&, int, int, cplx*, const FastTabs&);
static int g_fast = 0;  // set by emul_set_fast: run the compile-time-plan phase functions when a plan matches
extern "C" void emul_set_fast(int f) { g_fast = f; }
static void zinv_g(int ph, int tid, int nthr, const VlocTables& t, const VlocIO& io, int g, int b, cplx* s, const FastTabs&) { zinv_phase(ph, tid, nthr, t, io, g, b, s); }
static void zfwd_g(int ph, int tid, int nthr, const VlocTables& t, const VlocIO& io, int g, int b, cplx* s, const FastTabs&) { zfwd_phase(ph, tid, nthr, t, io, g, b, s); }
static void yinv_g(int ph, int tid, int nthr, const VlocTables& t, int z, int x, int b, cplx* s, const FastTabs&) { yinv_phase(ph, tid, nthr, t, z, x, b, s); }
static void yfwd_g(int ph, int tid, int nthr, const VlocTables& t, int z, int x, int b, cplx* s, const FastTabs&) { yfwd_phase(ph, tid, nthr, t, z, x, b, s); }
static void xfus_g(int ph, int tid, int nthr, const VlocTables& t, int tile, int b, cplx* s, const FastTabs&) { xfus_phase(ph, tid, nthr, t, tile, b, s); }
static zfn pick_z(const FftPlan& pl, bool inv, int = 32)
{
#define X(PL) if (g_fast && PL::matches(pl)) return inv ? (zfn)zinv_fast_phase<PL, 1> : (zfn)zfwd_fast_phase<PL, 1>;
    QEB_FOR_PLANS(X)
#undef X
    return inv ? zinv_g : zfwd_g;
}
// staged z-forward kernel form (vloc_fast.cuh::zfwd_stage + the STAGE phases): what the fast launcher runs unless QEB_ZSTAGE=0
static zfn pick_zfwd_staged(const FftPlan& pl)
{
    const char* e = getenv("QEB_ZSTAGE");
    if (!g_fast || (e && atoi(e) == 0)) return nullptr;
#define X(PL) if (PL::matches(pl)) return (zfn)zfwd_fast_phase<PL, 1, true>;
    QEB_FOR_PLANS(X)
#undef X
    return nullptr;
}
static void run_zfwd_group(const VlocTables& t, const VlocIO& io, int g, int box, cplx* smem, int nthr)
{
    const FastTabs ft{t.pz.tw, nullptr, t.pz.rev, t.grp_stick + (size_t)g * t.TZ};
    if (zfn f = pick_zfwd_staged(t.pz)) {
        std::vector<int> irev(t.nr3);
        for (int i = 0; i < t.nr3; ++i) irev[t.pz.rev[i]] = i;
        for (int tid = 0; tid < nthr; ++tid) zfwd_stage(tid, nthr, t, g, box, smem, irev.data(), ft.sticks);
        for (int ph = 0; ph < zfwd_nphase(t); ++ph)
            for (int tid = 0; tid < nthr; ++tid) f(ph, tid, nthr, t, io, g, box, smem, ft);
        return;
    }
    for (int ph = 0; ph < zfwd_nphase(t); ++ph)
        for (int tid = 0; tid < nthr; ++tid) pick_z(t.pz, false, t.TZ)(ph, tid, nthr, t, io, g, box, smem, ft);
}
static yfn pick_y(const FftPlan& pl, bool inv)
{
#define X(PL) if (g_fast && PL::matches(pl)) return inv ? (yfn)yinv_fast_phase<PL> : (yfn)yfwd_fast_phase<PL>;
    QEB_FOR_PLANS(X)
#undef X
    return inv ? yinv_g : yfwd_g;
}
static xfn pick_x(const FftPlan& pl)
{
#define X(PL) if (g_fast && PL::matches(pl)) return (xfn)xfus_fast_phase<PL>;
    QEB_FOR_PLANS(X)
#undef X
    return xfus_g;
}

// fused plane pass (vloc_plane.cuh): one plane z of one box, stick row in / out through g1
typedef void (*pfn)(int, int, int, const VlocTables&, int, cplx*, const cplx*, cplx*, const PlaneSm&);
#define QEB_FOR_PLANES(X) X(SP125, SP64) X(SP25, SP25) X(SP27, SP27) X(SP15, SP15) X(SP64, SP64) X(SP125b, SP64) X(SP125b, SP64b)
static pfn pick_plane(const VlocTables& t, int* nph)
{
#define X(PX, PY) if (PX::matches(t.px) && PY::matches(t.py)) { *nph = plane_nphase<PX, PY>(); return (pfn)plane_phase<PX, PY>; }
    QEB_FOR_PLANES(X)
#undef X
    return nullptr;
}

// second form of the plane pass (vloc_plane2.cuh): g_fast == 3. Same stick rows in / out; tables as the kernel stages them.
typedef void (*p2fn)(int, int, int, const VlocTables&, int, cplx*, const cplx*, cplx*, const Plane2Sm&);
typedef void (*p2mask)(const int*, int*);
#define QEB_FOR_PLANE2S(X) X(SP125, SP64) X(SP125b, SP64) X(SP25, SP25) X(SP27, SP27) X(SP15, SP15) X(SP64, SP64)
static p2fn pick_plane2(const VlocTables& t, int* nph, int* ldp, int* nmask, p2mask* mk)
{
#define X(PX, PY) if (PX::matches(t.px) && PY::matches(t.py)) { *nph = Plane2<PX, PY>::NPH; *ldp = Plane2<PX, PY>::LDP; *nmask = Plane2<PX, PY>::XMASK; *mk = (p2mask)plane2_xmask<PX>; return (p2fn)plane2_phase<PX, PY>; }
    QEB_FOR_PLANE2S(X)
#undef X
    return nullptr;
}

extern "C" int emul_vloc(int nr1, int nr2, int nr3, int npw, const int* nl, const int* nlm, const int* igk, int ngm,
                         int lda, int m, const double* psi, const double* v, const double* g2kin, double* hpsi,
                         int nthr, int batch)
{
    HostVlocTables H;
    std::string err = build_vloc_tables(nr1, nr2, nr3, npw, nl, nlm, igk, ngm, H);
    if (!err.empty()) {
        fprintf(stderr, "emul_vloc: %s\n", err.c_str());
        return 1;
    }
    VlocTables t;
    memset(&t, 0, sizeof(t));
    t.nr1 = nr1; t.nr2 = nr2; t.nr3 = nr3;
    t.nst = H.nst; t.nxw = H.nxw; t.npos = nr2 * nr3;
    upload_plan(H.px, t.px); upload_plan(H.py, t.py); upload_plan(H.pz, t.pz);
    t.xmap = H.xmap.data(); t.plane_s0 = H.plane_s0.data(); t.plane_ns = H.plane_ns.data(); t.ymap = H.ymap.data();
    t.TZ = H.TZ; t.ngrp = H.ngrp;
    t.grp_stick = H.grp_stick.data(); t.grp_nl = H.grp_nl.data(); t.grp_div = H.grp_div.data();
    t.grp_coef_off = H.grp_coef_off.data(); t.coef_ig = H.coef_ig.data(); t.coef_slot = H.coef_slot.data();
    t.coef_slot2 = H.gamma ? H.coef_slot2.data() : nullptr;
    t.TY = H.TY; t.nty = H.nty; t.divTY = make_fastdiv(H.TY);
    { int last = nr3 - (H.nty - 1) * H.TY; t.divTYlast = make_fastdiv(last); }
    t.TX = H.TX; t.ntx = H.ntx; t.divTX = make_fastdiv(H.TX);
    { int last = t.npos - (H.ntx - 1) * H.TX; t.divTXlast = make_fastdiv(last); }
    t.nsmax = H.nsmax;
    t.pl_ymapT = H.pl_ymapT.data(); t.pl_xact = H.plane_x.data(); t.pl_ldp = H.pl_ldp;
    std::vector<double> vpl((size_t)nr1 * nr2 * nr3);
    for (int z = 0; z < nr3; ++z)
        for (int xp = 0; xp < nr1; ++xp)
            for (int yp = 0; yp < nr2; ++yp)
                vpl[((size_t)z * nr1 + xp) * nr2 + yp] =
                    v[H.px.rev[xp] + (size_t)nr1 * (H.py.rev[yp] + (size_t)nr2 * z)] * (1.0 / ((double)nr1 * nr2 * nr3));
    t.pl_v = vpl.data();
    int plane_nph = 0;
    pfn plane_fn = g_fast == 2 ? pick_plane(t, &plane_nph) : nullptr;
    if (g_fast == 2 && !plane_fn) return 2;  // no instantiated plane kernel for this grid
    int p2_nph = 0, p2_ldp = 0, p2_nmask = 0;
    p2mask p2_mk = nullptr;
    p2fn plane2_fn = g_fast == 3 ? pick_plane2(t, &p2_nph, &p2_ldp, &p2_nmask, &p2_mk) : nullptr;
    if (g_fast == 3 && !plane2_fn) return 2;
    // potential: permuted by rev_x and scaled
    std::vector<double> vp((size_t)nr1 * t.npos);
    const double sc = 1.0 / ((double)nr1 * nr2 * nr3);
    for (int xp = 0; xp < nr1; ++xp) {
        int x = H.px.rev[xp];
        for (int y = 0; y < nr2; ++y)
            for (int z = 0; z < nr3; ++z)
                vp[(size_t)xp * t.npos + (size_t)y * nr3 + z] = v[x + (size_t)nr1 * (y + (size_t)nr2 * z)] * sc;
    }
    t.vp = vp.data();
    t.g1_box = (size_t)H.nst * nr3;
    t.g1_zs = (size_t)H.nst;
    t.g2_box = (size_t)H.nxw * t.npos;
    std::vector<cplx> g1(t.g1_box * batch), g2(t.g2_box * batch);
    t.g1 = g1.data(); t.g2 = g2.data();
    const cplx nanv = mk(0.0 / 0.0, 0.0 / 0.0);
    std::vector<cplx> smem(std::max(std::max(H.smem_z, H.smem_y), H.smem_x) / 16 + 16);

    VlocIO io;
    io.psi = (const cplx*)psi; io.hpsi = (cplx*)hpsi; io.g2kin = g2kin; io.lda = lda; io.n = npw; io.nbands = m;
    io.gamma = H.gamma;
    const int nboxes_total = H.gamma ? (m + 1) / 2 : m;
    for (int bx0 = 0; bx0 < nboxes_total; bx0 += batch) {
        const int nb = std::min(batch, nboxes_total - bx0);
        io.band0 = H.gamma ? 2 * bx0 : bx0;
        for (auto& x : g1) x = nanv;  // poison: every element read must have been written
        for (auto& x : g2) x = nanv;
        for (int box = 0; box < nb; ++box) {
            for (int g = 0; g < H.ngrp; ++g) {
                for (auto& x : smem) x = nanv;
                for (int ph = 0; ph < zinv_nphase(t); ++ph)
                    for (int tid = 0; tid < nthr; ++tid) pick_z(t.pz, true, t.TZ)(ph, tid, nthr, t, io, g, box, smem.data(), FastTabs{t.pz.tw, nullptr, t.pz.rev, t.grp_stick + (size_t)g * t.TZ});
            }
            if (plane2_fn) {
                std::vector<int> xmask(p2_nmask);
                p2_mk(t.xmap, xmask.data());
                std::vector<int> ymap2((size_t)nr2 * H.nxw);
                for (size_t i = 0; i < ymap2.size(); ++i) ymap2[i] = plane2_ymap(t.pl_ymapT[i], H.nst);
                const Plane2Sm psm{t.px.tw, t.py.tw, ymap2.data(), t.pl_xact, xmask.data()};
                std::vector<cplx> P((size_t)nr2 * p2_ldp), sin(plane2_row_elems(H.nst)), sout(plane2_row_elems(H.nst));
                const int pthr = ((nthr + 31) / 32) * 32;  // whole warps
                for (int z = 0; z < nr3; ++z) {
                    for (auto& x : P) x = nanv;
                    for (auto& x : sout) x = nanv;
                    cplx* row = t.g1 + (size_t)box * t.g1_box + (size_t)z * t.g1_zs;
                    for (int i = 0; i < H.nst; ++i) sin[i] = row[i];
                    sin[H.nst] = mk(0.0, 0.0);
                    for (int ph = 0; ph < p2_nph; ++ph)
                        for (int tid = 0; tid < pthr; ++tid) plane2_fn(ph, tid, pthr, t, z, P.data(), sin.data(), sout.data(), psm);
                    for (int i = 0; i < H.nst; ++i) row[i] = sout[i];
                }
            } else if (plane_fn) {
                const PlaneSm psm{t.px.tw, t.py.tw, t.pl_ymapT, t.xmap, t.pl_xact};
                std::vector<cplx> P((size_t)nr2 * t.pl_ldp), sin(H.nst), sout(H.nst);
                const int pthr = ((nthr + 31) / 32) * 32;  // whole warps
                for (int z = 0; z < nr3; ++z) {
                    for (auto& x : P) x = nanv;
                    for (auto& x : sout) x = nanv;
                    cplx* row = t.g1 + (size_t)box * t.g1_box + (size_t)z * t.g1_zs;
                    for (int i = 0; i < H.nst; ++i) sin[i] = row[i];
                    for (int ph = 0; ph < plane_nph; ++ph)
                        for (int tid = 0; tid < pthr; ++tid) plane_fn(ph, tid, pthr, t, z, P.data(), sin.data(), sout.data(), psm);
                    for (int i = 0; i < H.nst; ++i) row[i] = sout[i];
                }
            } else {
            for (int xa = 0; xa < H.nxw; ++xa)
                for (int zt = 0; zt < H.nty; ++zt) {
                    for (auto& x : smem) x = nanv;
                    for (int ph = 0; ph < ypass_nphase(t); ++ph)
                        for (int tid = 0; tid < nthr; ++tid) pick_y(t.py, true)(ph, tid, nthr, t, zt, xa, box, smem.data(), FastTabs{t.py.tw, t.ymap + (size_t)xa * t.nr2, t.py.rev, nullptr});
                }
            for (int tile = 0; tile < H.ntx; ++tile) {
                for (auto& x : smem) x = nanv;
                for (int ph = 0; ph < xfus_nphase(t); ++ph)
                    for (int tid = 0; tid < nthr; ++tid) pick_x(t.px)(ph, tid, nthr, t, tile, box, smem.data(), FastTabs{t.px.tw, t.xmap, nullptr, nullptr});
            }
            for (int xa = 0; xa < H.nxw; ++xa)
                for (int zt = 0; zt < H.nty; ++zt) {
                    for (auto& x : smem) x = nanv;
                    for (int ph = 0; ph < ypass_nphase(t); ++ph)
                        for (int tid = 0; tid < nthr; ++tid) pick_y(t.py, false)(ph, tid, nthr, t, zt, xa, box, smem.data(), FastTabs{t.py.tw, t.ymap + (size_t)xa * t.nr2, t.py.rev, nullptr});
                }
            }
            for (int g = 0; g < H.ngrp; ++g) {
                for (auto& x : smem) x = nanv;
                run_zfwd_group(t, io, g, box, smem.data(), nthr);
            }
        }
    }
    return 0;
}

// sum_band density loop: zinv + yinv phases as in emul_vloc, then the xacc phases (generic, or the static-plan variant
// when g_fast is set) with the accumulator flushed into rho once per batch. rho (nnr, x fastest) is INOUT.
extern "C" int emul_sumband(int nr1, int nr2, int nr3, int npw, const int* nl, const int* nlm, const int* igk, int ngm,
                            int lda, int m, const double* psi, const double* w, double* rho, int nthr, int batch)
{
    HostVlocTables H;
    std::string err = build_vloc_tables(nr1, nr2, nr3, npw, nl, nlm, igk, ngm, H);
    if (!err.empty()) return 1;
    VlocTables t;
    memset(&t, 0, sizeof(t));
    t.nr1 = nr1; t.nr2 = nr2; t.nr3 = nr3;
    t.nst = H.nst; t.nxw = H.nxw; t.npos = nr2 * nr3;
    upload_plan(H.px, t.px); upload_plan(H.py, t.py); upload_plan(H.pz, t.pz);
    t.xmap = H.xmap.data(); t.plane_s0 = H.plane_s0.data(); t.plane_ns = H.plane_ns.data(); t.ymap = H.ymap.data();
    t.TZ = H.TZ; t.ngrp = H.ngrp;
    t.grp_stick = H.grp_stick.data(); t.grp_nl = H.grp_nl.data(); t.grp_div = H.grp_div.data();
    t.grp_coef_off = H.grp_coef_off.data(); t.coef_ig = H.coef_ig.data(); t.coef_slot = H.coef_slot.data();
    t.coef_slot2 = H.gamma ? H.coef_slot2.data() : nullptr;
    t.TY = H.TY; t.nty = H.nty; t.divTY = make_fastdiv(H.TY);
    { int last = nr3 - (H.nty - 1) * H.TY; t.divTYlast = make_fastdiv(last); }
    t.TX = H.TX; t.ntx = H.ntx; t.divTX = make_fastdiv(H.TX);
    { int last = t.npos - (H.ntx - 1) * H.TX; t.divTXlast = make_fastdiv(last); }
    t.nsmax = H.nsmax;
    t.g1_box = (size_t)H.nst * nr3; t.g1_zs = (size_t)H.nst; t.g2_box = (size_t)H.nxw * t.npos;
    std::vector<cplx> g1(t.g1_box * batch), g2(t.g2_box * batch);
    t.g1 = g1.data(); t.g2 = g2.data();
    const cplx nanv = mk(0.0 / 0.0, 0.0 / 0.0);
    std::vector<cplx> smem(std::max(std::max(H.smem_z, H.smem_y), H.smem_x) / 16 + 16);
    std::vector<double> racc((size_t)nr1 * H.TX);
    std::vector<int> irev(nr1);
    for (int i = 0; i < nr1; ++i) irev[H.px.rev[i]] = i;
    VlocIO io;
    io.psi = (const cplx*)psi; io.hpsi = nullptr; io.g2kin = nullptr; io.lda = lda; io.n = npw; io.nbands = m; io.gamma = H.gamma;
    const int nboxes_total = H.gamma ? (m + 1) / 2 : m;
    const int pthr = ((nthr + 31) / 32) * 32;
    for (int bx0 = 0; bx0 < nboxes_total; bx0 += batch) {
        const int nb = std::min(batch, nboxes_total - bx0);
        io.band0 = H.gamma ? 2 * bx0 : bx0;
        for (auto& x : g1) x = nanv;
        for (auto& x : g2) x = nanv;
        for (int box = 0; box < nb; ++box) {
            for (int g = 0; g < H.ngrp; ++g) {
                for (auto& x : smem) x = nanv;
                for (int ph = 0; ph < zinv_nphase(t); ++ph)
                    for (int tid = 0; tid < nthr; ++tid) pick_z(t.pz, true, t.TZ)(ph, tid, nthr, t, io, g, box, smem.data(), FastTabs{t.pz.tw, nullptr, t.pz.rev, t.grp_stick + (size_t)g * t.TZ});
            }
            for (int xa = 0; xa < H.nxw; ++xa)
                for (int zt = 0; zt < H.nty; ++zt) {
                    for (auto& x : smem) x = nanv;
                    for (int ph = 0; ph < ypass_nphase(t); ++ph)
                        for (int tid = 0; tid < nthr; ++tid) pick_y(t.py, true)(ph, tid, nthr, t, zt, xa, box, smem.data(), FastTabs{t.py.tw, t.ymap + (size_t)xa * t.nr2, t.py.rev, nullptr});
                }
        }
        for (int tile = 0; tile < H.ntx; ++tile) {
            for (int tid = 0; tid < pthr; ++tid) xacc_zero_phase(tid, pthr, t, racc.data());
            for (int box = 0; box < nb; ++box) {
                for (auto& x : smem) x = nanv;
                for (int ph = 0; ph < t.px.npass; ++ph)
                    for (int tid = 0; tid < pthr; ++tid) {
                        bool done = false;
                        if (g_fast) {
#define X(PL) if (!done && PL::matches(t.px)) { xacc_fast_phase<PL>(ph, tid, pthr, t, io, w, tile, box, smem.data(), racc.data(), FastTabs{t.px.tw, t.xmap, nullptr, nullptr}); done = true; }
                            QEB_FOR_PLANS(X)
#undef X
                        }
                        if (!done) xacc_phase(ph, tid, pthr, t, io, w, tile, box, smem.data(), racc.data());
                    }
            }
            for (int tid = 0; tid < pthr; ++tid) xacc_flush_phase(tid, pthr, t, tile, irev.data(), racc.data(), rho);
        }
    }
    return 0;
}

// noncollinear V_loc with the 2x2 potential: zinv / yinv phases on the 2m component boxes, the spinor x pass
// (xspin_phase) per band, yfwd / zfwd. psi, hpsi: 2m columns of lda; v4: (4, nnr) x fastest. hpsi INOUT.
extern "C" int emul_vloc_nc(int nr1, int nr2, int nr3, int npw, const int* nl, const int* igk, int ngm, int lda, int m,
                            const double* psi, const double* v4, double* hpsi, int nthr, int batch_bands)
{
    HostVlocTables H;
    std::string err = build_vloc_tables(nr1, nr2, nr3, npw, nl, nullptr, igk, ngm, H);
    if (!err.empty()) return 1;
    VlocTables t;
    memset(&t, 0, sizeof(t));
    t.nr1 = nr1; t.nr2 = nr2; t.nr3 = nr3;
    t.nst = H.nst; t.nxw = H.nxw; t.npos = nr2 * nr3;
    upload_plan(H.px, t.px); upload_plan(H.py, t.py); upload_plan(H.pz, t.pz);
    t.xmap = H.xmap.data(); t.plane_s0 = H.plane_s0.data(); t.plane_ns = H.plane_ns.data(); t.ymap = H.ymap.data();
    t.TZ = H.TZ; t.ngrp = H.ngrp;
    t.grp_stick = H.grp_stick.data(); t.grp_nl = H.grp_nl.data(); t.grp_div = H.grp_div.data();
    t.grp_coef_off = H.grp_coef_off.data(); t.coef_ig = H.coef_ig.data(); t.coef_slot = H.coef_slot.data();
    t.TY = H.TY; t.nty = H.nty; t.divTY = make_fastdiv(H.TY);
    { int last = nr3 - (H.nty - 1) * H.TY; t.divTYlast = make_fastdiv(last); }
    t.TX = H.TX; t.ntx = H.ntx; t.divTX = make_fastdiv(H.TX);
    { int last = t.npos - (H.ntx - 1) * H.TX; t.divTXlast = make_fastdiv(last); }
    t.nsmax = H.nsmax;
    const int batch = 2 * batch_bands;
    t.g1_box = (size_t)H.nst * nr3; t.g1_zs = (size_t)H.nst; t.g2_box = (size_t)H.nxw * t.npos;
    std::vector<cplx> g1(t.g1_box * batch), g2(t.g2_box * batch);
    t.g1 = g1.data(); t.g2 = g2.data();
    const size_t nnr = (size_t)nr1 * nr2 * nr3;
    const double scale = 1.0 / (double)nnr;
    std::vector<double> vnc(4 * nnr);
    for (int k = 0; k < 4; ++k)
        for (int xp = 0; xp < nr1; ++xp)
            for (int y = 0; y < nr2; ++y)
                for (int z = 0; z < nr3; ++z)
                    vnc[k * nnr + (size_t)xp * t.npos + (size_t)y * nr3 + z] =
                        v4[k * nnr + H.px.rev[xp] + (size_t)nr1 * (y + (size_t)nr2 * z)] * scale;
    t.vnc = vnc.data();
    const cplx nanv = mk(0.0 / 0.0, 0.0 / 0.0);
    std::vector<cplx> smem(std::max(std::max(H.smem_z, H.smem_y), 2 * H.smem_x) / 16 + 16);
    VlocIO io;
    io.psi = (const cplx*)psi; io.hpsi = (cplx*)hpsi; io.g2kin = nullptr; io.lda = lda; io.n = npw; io.nbands = 2 * m; io.gamma = 0;
    for (int b0 = 0; b0 < m; b0 += batch_bands) {
        const int nbands = std::min(batch_bands, m - b0);
        const int nb = 2 * nbands;
        io.band0 = 2 * b0;
        for (auto& x : g1) x = nanv;
        for (auto& x : g2) x = nanv;
        for (int box = 0; box < nb; ++box) {
            for (int g = 0; g < H.ngrp; ++g) {
                for (auto& x : smem) x = nanv;
                for (int ph = 0; ph < zinv_nphase(t); ++ph)
                    for (int tid = 0; tid < nthr; ++tid) pick_z(t.pz, true, t.TZ)(ph, tid, nthr, t, io, g, box, smem.data(), FastTabs{t.pz.tw, nullptr, t.pz.rev, t.grp_stick + (size_t)g * t.TZ});
            }
            for (int xa = 0; xa < H.nxw; ++xa)
                for (int zt = 0; zt < H.nty; ++zt) {
                    for (auto& x : smem) x = nanv;
                    for (int ph = 0; ph < ypass_nphase(t); ++ph)
                        for (int tid = 0; tid < nthr; ++tid) pick_y(t.py, true)(ph, tid, nthr, t, zt, xa, box, smem.data(), FastTabs{t.py.tw, t.ymap + (size_t)xa * t.nr2, t.py.rev, nullptr});
                }
        }
        for (int pair = 0; pair < nbands; ++pair)
            for (int tile = 0; tile < H.ntx; ++tile) {
                for (auto& x : smem) x = nanv;
                bool done = false;
                if (g_fast) {
#define X(PL) if (!done && PL::matches(t.px)) { \
                    for (int ph = 0; ph < 2 * PL::NP - 1; ++ph) \
                        for (int tid = 0; tid < nthr; ++tid) xspin_fast_phase<PL>(ph, tid, nthr, t, tile, pair, smem.data(), FastTabs{t.px.tw, t.xmap, nullptr, nullptr}); \
                    done = true; }
                    QEB_FOR_PLANS(X)
#undef X
                }
                if (!done)
                    for (int ph = 0; ph < xspin_nphase(t); ++ph)
                        for (int tid = 0; tid < nthr; ++tid) xspin_phase(ph, tid, nthr, t, tile, pair, smem.data());
            }
        for (int box = 0; box < nb; ++box) {
            for (int xa = 0; xa < H.nxw; ++xa)
                for (int zt = 0; zt < H.nty; ++zt) {
                    for (auto& x : smem) x = nanv;
                    for (int ph = 0; ph < ypass_nphase(t); ++ph)
                        for (int tid = 0; tid < nthr; ++tid) pick_y(t.py, false)(ph, tid, nthr, t, zt, xa, box, smem.data(), FastTabs{t.py.tw, t.ymap + (size_t)xa * t.nr2, t.py.rev, nullptr});
                }
            for (int g = 0; g < H.ngrp; ++g) {
                for (auto& x : smem) x = nanv;
                run_zfwd_group(t, io, g, box, smem.data(), nthr);
            }
        }
    }
    return 0;
}

extern "C" int emul_fullfft(int nr1, int nr2, int nr3, int howmany, int sign, double* f, int nthr)
{
    HostPlan hx, hy, hz;
    if (!make_host_plan(nr1, hx) || !make_host_plan(nr2, hy) || !make_host_plan(nr3, hz)) return 1;
    FullFftParams p;
    p.nr1 = nr1; p.nr2 = nr2; p.nr3 = nr3;
    upload_plan(hx, p.px); upload_plan(hy, p.py); upload_plan(hz, p.pz);
    p.f = (cplx*)f;
    p.sgn = sign > 0 ? 1.0 : -1.0;
    p.scale = sign > 0 ? 1.0 : 1.0 / ((double)nr1 * nr2 * nr3);
    const int ncol = nr1 * nr2, nrows = nr2 * nr3;
    p.TZ = std::min(6, ncol); p.TYx = std::min(4, nr1); p.TXr = std::min(5, nrows);  // small odd tiles: exercise the edges
    const int ntz = (ncol + p.TZ - 1) / p.TZ, nty = (nr1 + p.TYx - 1) / p.TYx, ntx = (nrows + p.TXr - 1) / p.TXr;
    p.dTZ = make_fastdiv(p.TZ); p.dTZl = make_fastdiv(ncol - (ntz - 1) * p.TZ);
    p.dTY = make_fastdiv(p.TYx); p.dTYl = make_fastdiv(nr1 - (nty - 1) * p.TYx);
    p.dTX = make_fastdiv(p.TXr); p.dTXl = make_fastdiv(nrows - (ntx - 1) * p.TXr);
    std::vector<cplx> smem((size_t)std::max(std::max(nr3 * p.TZ, nr2 * p.TYx), nr1 * (p.TXr + 1)) + 16);
    for (int box = 0; box < howmany; ++box) {
        for (int t = 0; t < ntz; ++t)
            for (int ph = 0; ph < p.pz.npass; ++ph)
                for (int tid = 0; tid < nthr; ++tid) fullz_phase(ph, tid, nthr, p, t, box, smem.data());
        for (int z = 0; z < nr3; ++z)
            for (int t = 0; t < nty; ++t)
                for (int ph = 0; ph < p.py.npass; ++ph)
                    for (int tid = 0; tid < nthr; ++tid) fully_phase(ph, tid, nthr, p, t, z, box, smem.data());
        for (int t = 0; t < ntx; ++t)
            for (int ph = 0; ph < fullx_nphase(p); ++ph)
                for (int tid = 0; tid < nthr; ++tid) fullx_phase(ph, tid, nthr, p, t, box, smem.data());
    }
    return 0;
}
