// Host-side construction of the stick / plane / coefficient tables the V_loc
// kernels consume, from the reference's own index maps:
//   nl, nlm  (fft_set_nl, FFTXlib/src/fft_ggen.f90:28-108; serial layout
//             nl = 1 + n1 + n2*nr1x + n3*nr1x*nr2x, :79)
//   igk      (PW/src/gk_sort.f90:9-98; igk_k(:,current_k), PW/src/pwcom.f90:72-105)
// The reference prunes its 3-D transform with dfft%isind / dfft%iplw, which
// cover the union of all k-points (fft_types.f90:814-851); we prune with the
// exact stick set of the current k-point, a subset: columns that hold no
// coefficient transform to zero (G->R) and are never gathered (R->G), so the
// results are identical.
#pragma once
#include <cstdlib>
#include <algorithm>
#include <string>
#include <vector>
#include "fft_plan.h"
#include "vloc_kernels.cuh"
#include "vloc_plane.cuh"

namespace qeb {

struct HostVlocTables {
    int nr1 = 0, nr2 = 0, nr3 = 0, nst = 0, nxw = 0, nsmax = 0;
    int TZ = 0, TY = 0, TX = 0, ngrp = 0, nty = 0, ntx = 0;
    int gamma = 0;
    HostPlan px, py, pz;
    std::vector<int> xmap, plane_x, plane_s0, plane_ns, ymap;
    std::vector<int> stick_x, stick_y;
    std::vector<int> grp_stick, grp_nl, grp_coef_off, coef_ig, coef_slot, coef_slot2;
    std::vector<FastDiv> grp_div;
    size_t smem_z = 0, smem_y = 0, smem_x = 0;
    // fused plane pass (vloc_plane.cuh)
    std::vector<int> pl_ymapT;  // [nr2*nxw] stick id at (y, xa) or -1
    int pl_ldp = 0;
    size_t smem_plane = 0;
};

inline int pick_tile(int n, int lo, int hi)
{
    // tile length in [lo,hi] (even when possible) with the least padding
    if (n <= hi) return n;
    int best = hi;
    double beff = 0;
    for (int t = hi; t >= lo; --t) {
        if (t & 1) continue;
        int nt = (n + t - 1) / t;
        double eff = (double)n / ((double)nt * t);
        if (eff > beff + 1e-9) {
            beff = eff;
            best = t;
        }
    }
    return best;
}

// nl/nlm/igk are the reference's 1-based arrays. igk == nullptr => igk(i) = i.
// force_cols: box columns (x + nr1*y) that are sticks even if no coefficient lies on them (the slab-distributed
// path transforms every wave stick of the descriptor, because all ranks must agree on the stick lists).
// need_pz = false builds plane tables only (nr3 = number of local z planes, any value).
inline std::string build_vloc_tables(int nr1, int nr2, int nr3, int npw, const int* nl, const int* nlm,
                                     const int* igk, int ngm, HostVlocTables& T, size_t smem_cap = 72 * 1024,
                                     const int* force_cols = nullptr, int nforce = 0, bool need_pz = true,
                                     bool gamma_flag = false)
{
    T = HostVlocTables();
    T.nr1 = nr1;
    T.nr2 = nr2;
    T.nr3 = nr3;
    T.gamma = nlm != nullptr || gamma_flag;
    if (!make_host_plan(nr1, T.px, 10, getenv("QEB_XPLAN")) || !make_host_plan(nr2, T.py, 10, getenv("QEB_YPLAN")) ||
        (need_pz && !make_host_plan(nr3, T.pz, 10, getenv("QEB_ZPLAN"))))
        return "FFT dimension with a prime factor other than 2, 3, 5";
    const int ncol = nr1 * nr2;
    std::vector<int> col_stick(ncol, -1);
    std::vector<int> cpos(npw), cneg(T.gamma ? npw : 0);
    for (int i = 0; i < npw; ++i) {
        int ig = igk ? igk[i] : (i + 1);
        if (ig < 1 || ig > ngm) return "igk out of range";
        int b = nl[ig - 1] - 1;
        if (b < 0 || b >= ncol * nr3) return "nl out of range";
        cpos[i] = b;
        col_stick[b % ncol] = -2;
        if (nlm) {
            int bm = nlm[ig - 1] - 1;
            if (bm < 0 || bm >= ncol * nr3) return "nlm out of range";
            cneg[i] = bm;
            col_stick[bm % ncol] = -2;
        }
    }
    for (int i = 0; i < nforce; ++i) {
        if (force_cols[i] < 0 || force_cols[i] >= ncol) return "stick column out of range";
        col_stick[force_cols[i]] = -2;
    }
    // sticks ordered by (x, y); planes by x
    T.xmap.assign(nr1, -1);
    for (int x = 0; x < nr1; ++x) {
        int ns = 0;
        for (int y = 0; y < nr2; ++y)
            if (col_stick[x + nr1 * y] == -2) ++ns;
        if (!ns) continue;
        T.xmap[x] = T.nxw++;
        T.plane_x.push_back(x);
        T.plane_s0.push_back(T.nst);
        T.plane_ns.push_back(ns);
        T.nsmax = std::max(T.nsmax, ns);
        int l = 0;
        for (int y = 0; y < nr2; ++y)
            if (col_stick[x + nr1 * y] == -2) {
                col_stick[x + nr1 * y] = T.nst++;
                T.stick_x.push_back(x);
                T.stick_y.push_back(y);
                ++l;
            }
    }
    T.ymap.assign((size_t)T.nxw * nr2, -1);
    for (int s = 0; s < T.nst; ++s) {
        int xa = T.xmap[T.stick_x[s]];
        T.ymap[(size_t)xa * nr2 + T.stick_y[s]] = s - T.plane_s0[xa];
    }
    // fused plane pass: transposed stick map, plane row stride (odd)
    T.pl_ymapT.assign((size_t)nr2 * T.nxw, -1);
    for (int s = 0; s < T.nst; ++s) T.pl_ymapT[(size_t)T.stick_y[s] * T.nxw + T.xmap[T.stick_x[s]]] = s;
    T.pl_ldp = nr1 | 1;
    T.smem_plane = plane_smem_bytes(nr1, nr2, T.nst, T.nxw, T.pl_ldp);
    // tile sizes
    {
        int cap = (int)(smem_cap / (16 * (size_t)nr3));
        int tz = std::min(32, std::max(1, cap));
        if (tz >= 8) tz &= ~7;
        if (const char* e = getenv("QEB_TZ")) tz = std::max(1, std::min(32, atoi(e)));  // tuning aid (tools/vloc_sweep.py)
        if (T.gamma && tz < 2) tz = 2;
        T.TZ = std::max(1, std::min(tz, std::max(T.nst, T.gamma ? 2 : 1)));
        if (T.gamma && T.TZ < 2) T.TZ = 2;
        T.smem_z = (size_t)nr3 * T.TZ * 16;
        // y tiles: at most 32 z-lines (one per lane in the fast kernels), evenly sized
        int ty_hi = 32;
        while (ty_hi > 8 && ((size_t)nr2 * ty_hi + (size_t)T.nsmax * (ty_hi + 1)) * 16 > smem_cap) ty_hi -= 2;
        {
            const int nt = (nr3 + ty_hi - 1) / ty_hi;
            T.TY = (nr3 + nt - 1) / nt;
            // full 32-line tiles keep every 512-byte G2 row segment sector-aligned; measured +2% on the V_loc section
            // at nr3 = 200 (profiles/r1m_sweep.log) - taken when the short last tile wastes little
            if (ty_hi == 32 && (nt * 32 - nr3) * 100 <= 12 * nt * 32) T.TY = 32;
        }
        if (const char* e = getenv("QEB_TY")) T.TY = std::max(1, std::min(ty_hi, atoi(e)));  // tuning aid
        T.nty = (nr3 + T.TY - 1) / T.TY;
        T.smem_y = ((size_t)nr2 * T.TY + (size_t)T.nsmax * (T.TY + 1)) * 16;
        int tx_hi = 32;
        while (tx_hi > 4 && (size_t)nr1 * tx_hi * 16 > smem_cap) tx_hi -= 2;
        T.TX = std::min(tx_hi, nr2 * nr3);
        T.ntx = (nr2 * nr3 + T.TX - 1) / T.TX;
        T.smem_x = (size_t)nr1 * T.TX * 16;
    }
    // z-pass groups
    std::vector<int> stick_grp(T.nst, -1), stick_lane(T.nst, -1);
    {
        int cur = -1, fill = T.TZ;  // force a new group first
        auto new_group = [&]() {
            ++cur;
            fill = 0;
            T.grp_stick.insert(T.grp_stick.end(), T.TZ, -1);
            T.grp_nl.push_back(0);
        };
        auto add = [&](int s) {
            stick_grp[s] = cur;
            stick_lane[s] = fill;
            T.grp_stick[(size_t)cur * T.TZ + fill] = s;
            ++fill;
            T.grp_nl[cur] = fill;
        };
        for (int s = 0; s < T.nst; ++s) {
            if (stick_grp[s] >= 0) continue;
            int mirror = -1;
            if (T.gamma) {
                int mx = (nr1 - T.stick_x[s]) % nr1, my = (nr2 - T.stick_y[s]) % nr2;
                mirror = col_stick[mx + nr1 * my];
                if (mirror < 0) return "Gamma: mirror stick missing";
                if (mirror == s || stick_grp[mirror] >= 0) mirror = -1;
            }
            int need = mirror >= 0 ? 2 : 1;
            if (fill + need > T.TZ) new_group();
            add(s);
            if (mirror >= 0) add(mirror);
        }
        T.ngrp = cur + 1;
        for (int g = 0; g < T.ngrp; ++g) T.grp_div.push_back(make_fastdiv(T.grp_nl[g]));
    }
    // coefficient lists per group
    {
        std::vector<int> cnt(T.ngrp + 1, 0);
        std::vector<int> cg(npw);
        for (int i = 0; i < npw; ++i) {
            int s = col_stick[cpos[i] % ncol];
            cg[i] = stick_grp[s];
            if (T.gamma) {
                int sm = col_stick[cneg[i] % ncol];
                if (stick_grp[sm] != cg[i]) return "Gamma: +G/-G sticks in different groups";
            }
            ++cnt[cg[i] + 1];
        }
        for (int g = 0; g < T.ngrp; ++g) cnt[g + 1] += cnt[g];
        T.grp_coef_off = cnt;
        T.coef_ig.assign(npw, 0);
        T.coef_slot.assign(npw, 0);
        if (T.gamma) T.coef_slot2.assign(npw, 0);
        std::vector<int> fillc(cnt.begin(), cnt.end() - 1);
        for (int i = 0; i < npw; ++i) {
            int c = fillc[cg[i]]++;
            T.coef_ig[c] = i;
            int s = col_stick[cpos[i] % ncol];
            T.coef_slot[c] = (cpos[i] / ncol) * T.TZ + stick_lane[s];
            if (T.gamma) {
                int sm = col_stick[cneg[i] % ncol];
                T.coef_slot2[c] = (cneg[i] / ncol) * T.TZ + stick_lane[sm];
            }
        }
    }
    return "";
}

}  // namespace qeb
