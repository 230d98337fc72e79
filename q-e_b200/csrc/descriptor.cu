// Host-side descriptor builder of the library (no device code): the data the reference prepares once per run /
// per k-point and that defines the layouts the H|psi> kernels honour at the boundary:
//   grid dimensions        FFTXlib/src/fft_types.f90:1086-1122 (realspace_grid_init), :1128-1197 (grid_set),
//                          FFTXlib/src/fft_support.f90:61-165 (allowed / good_fft_order: factors 2, 3, 5)
//   G-vectors              Modules/recvec_subs.f90:111-258 (ggen: loop order, |G|^2 sort with eps8 ties,
//                          Gamma half-space rule, gstart), Modules/sort.f90:9-134 (hpsort_eps)
//   nl / nlm               FFTXlib/src/fft_ggen.f90:28-108 (fft_set_nl; serial x-fastest, parallel stick-major)
//   igk / g2kin            PW/src/gk_sort.f90:9-98, PW/src/g2_kin.f90:37-43
//   wave sticks / planes   FFTXlib/src/fft_types.f90:814-851 (serial isind / iplw)
//   slab distribution      FFTXlib/src/stick_base.f90:214-601 (sticks_map_set/index/sort/dist, get_sticks),
//                          FFTXlib/src/fft_types.f90:491-943 (fft_type_set: nr3p, iss, ismap, nsw, nsp)
// A Fortran host already owns these arrays and hands them to qeb_set_grid / qeb_set_grid_dist; hosts without the
// reference's setup code (the benchmark, the Python mirror, tests) get them from here.
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <map>
#include <string>
#include <vector>
#include "../../include/qeb200.h"

namespace {

constexpr double EPS8 = 1.0e-8;

// ascending heap sort of ra carrying ind; keys closer than eps are "equal" and are then ordered by ind
// (Modules/sort.f90:9-134). ind[0] == 0 on entry: initialise ind to 1..n.
void heap_sort_eps(int n, double* ra, int* ind, double eps)
{
    if (n > 0 && ind[0] == 0)
        for (int i = 0; i < n; ++i) ind[i] = i + 1;
    if (n < 2) return;
    auto before = [&](double a, int ia, double b, int ib) {  // (a,ia) sorts before (b,ib)
        if (fabs(a - b) >= eps) return a < b;
        return ia < ib;
    };
    int l = n / 2 + 1, ir = n;
    double rra;
    int iind;
    for (;;) {
        if (l > 1) {
            --l;
            rra = ra[l - 1];
            iind = ind[l - 1];
        } else {
            rra = ra[ir - 1];
            iind = ind[ir - 1];
            ra[ir - 1] = ra[0];
            ind[ir - 1] = ind[0];
            if (--ir == 1) {
                ra[0] = rra;
                ind[0] = iind;
                return;
            }
        }
        int i = l, j = l + l;
        while (j <= ir) {
            if (j < ir && before(ra[j - 1], ind[j - 1], ra[j], ind[j])) ++j;
            if (before(rra, iind, ra[j - 1], ind[j - 1])) {
                ra[i - 1] = ra[j - 1];
                ind[i - 1] = ind[j - 1];
                i = j;
                j += j;
            } else {
                j = ir + 1;
            }
        }
        ra[i - 1] = rra;
        ind[i - 1] = iind;
    }
}

bool fft_allowed(int nr)
{
    int mr = nr;
    const int fac[5] = {2, 3, 5, 7, 11};
    for (int f = 0; f < 5; ++f) {
        int pw = 0;
        while (mr % fac[f] == 0 && mr > 1) {
            mr /= fac[f];
            ++pw;
        }
        if (f >= 3 && pw > 0) return false;  // non-ESSL builds: 7 and 11 are not allowed
    }
    return mr == 1;
}

int good_order(int nr)
{
    int n = nr;
    while (!fft_allowed(n)) {
        if (++n > 16385) return -1;
    }
    return n;
}

struct RankData {
    std::vector<int> ig_l2g, nl, nlm, igk, pw_l2g;
    std::vector<double> g2kin;
    int gstart = 1;
};

}  // namespace

struct qeb_desc {
    double at[3][3], bg[3][3];  // [component][vector]
    double alat, tpiba2, gcutw, gcutms, gkcut, xk[3];
    bool gamma;
    int nr1, nr2, nr3, ngm, npw, nst, gstart;
    std::vector<int> mill, nl, nlm, igk, isind, iplw;  // mill: ngm x 3
    std::vector<double> g, gg, g2kin;                   // g: ngm x 3
    // distribution
    int nproc = 0;
    std::vector<int> nr3p, nsw, nsp, iss, ismap;
    std::vector<RankData> ranks;
    std::string err;
};

namespace {

void recips(qeb_desc* d)
{
    double a[3][3];  // a[v][c]
    for (int v = 0; v < 3; ++v)
        for (int c = 0; c < 3; ++c) a[v][c] = d->at[c][v];
    auto cross = [](const double* x, const double* y, double* o) {
        o[0] = x[1] * y[2] - x[2] * y[1];
        o[1] = x[2] * y[0] - x[0] * y[2];
        o[2] = x[0] * y[1] - x[1] * y[0];
    };
    double c23[3], c31[3], c12[3];
    cross(a[1], a[2], c23);
    cross(a[2], a[0], c31);
    cross(a[0], a[1], c12);
    const double den = a[0][0] * c23[0] + a[0][1] * c23[1] + a[0][2] * c23[2];
    for (int c = 0; c < 3; ++c) {
        d->bg[c][0] = c23[c] / den;
        d->bg[c][1] = c31[c] / den;
        d->bg[c][2] = c12[c] / den;
    }
}

inline double gnorm2(const qeb_desc* d, int i, int j, int k)
{
    double t[3];
    for (int c = 0; c < 3; ++c) t[c] = (i * d->bg[c][0] + j * d->bg[c][1]) + k * d->bg[c][2];
    return t[0] * t[0] + t[1] * t[1] + t[2] * t[2];
}

// number of G with |G|^2 <= gcut in column (i1,i2); Gamma: half space (stick_base.f90:214-262)
void column_counts(const qeb_desc* d, double gcut, int ub[3], std::vector<int>& st)
{
    ub[0] = (d->nr1 - 1) / 2;
    ub[1] = (d->nr2 - 1) / 2;
    ub[2] = (d->nr3 - 1) / 2;
    const int n1 = 2 * ub[0] + 1, n2 = 2 * ub[1] + 1;
    st.assign((size_t)n1 * n2, 0);
    for (int i1 = -ub[0]; i1 <= ub[0]; ++i1) {
        if (d->gamma && i1 < 0) continue;
        for (int i2 = -ub[1]; i2 <= ub[1]; ++i2) {
            if (d->gamma && i1 == 0 && i2 < 0) continue;
            int cnt = 0;
            for (int i3 = -ub[2]; i3 <= ub[2]; ++i3) {
                if (d->gamma && i1 == 0 && i2 == 0 && i3 < 0) continue;
                // same expression as the reference: sum over components of (i1 b1 + i2 b2 + i3 b3)^2
                double amod = 0.0;
                for (int c = 0; c < 3; ++c) {
                    const double t = i1 * d->bg[c][0] + i2 * d->bg[c][1] + i3 * d->bg[c][2];
                    amod += t * t;
                }
                if (amod <= gcut) ++cnt;
            }
            st[(size_t)(i1 + ub[0]) * n2 + (i2 + ub[1])] = cnt;
        }
    }
}

}  // namespace

namespace {

// realspace_grid_init + grid_set: nr_i = good_fft_order(2 nb_i + 1), nb_i = max |Miller index| with |G|^2 < gcut
int grid_dims(qeb_desc* d, double gcut)
{
    int est[3];
    for (int v = 0; v < 3; ++v) {
        const double len2 = d->at[0][v] * d->at[0][v] + d->at[1][v] * d->at[1][v] + d->at[2][v] * d->at[2][v];
        est[v] = (int)(sqrt(gcut) * sqrt(len2)) + 1;
    }
    int nb[3] = {0, 0, 0};
    for (int i = -est[0]; i <= est[0]; ++i)
        for (int j = -est[1]; j <= est[1]; ++j)
            for (int k = -est[2]; k <= est[2]; ++k) {
                double g[3];
                for (int c = 0; c < 3; ++c) g[c] = (i * d->bg[c][0] + j * d->bg[c][1]) + k * d->bg[c][2];
                const double gsq = (g[0] * g[0] + g[1] * g[1]) + g[2] * g[2];
                if (gsq < gcut) {  // strict: fft_types.f90:1171
                    nb[0] = std::max(nb[0], abs(i));
                    nb[1] = std::max(nb[1], abs(j));
                    nb[2] = std::max(nb[2], abs(k));
                }
            }
    d->nr1 = good_order(2 * nb[0] + 1);
    d->nr2 = good_order(2 * nb[1] + 1);
    d->nr3 = good_order(2 * nb[2] + 1);
    return (d->nr1 < 0 || d->nr2 < 0 || d->nr3 < 0) ? 1 : 0;
}

inline int wrap(int i, int n) { return i < 0 ? i + n : i; }

// ggen with global sort + fft_set_nl (serial layout)
void gen_gvectors(qeb_desc* d)
{
    const int ni = (d->nr1 - 1) / 2, nj = (d->nr2 - 1) / 2, nk = (d->nr3 - 1) / 2;
    std::vector<int> mu;       // unsorted Miller indices
    std::vector<double> g2;    // sort keys
    for (int i = d->gamma ? 0 : -ni; i <= ni; ++i) {
        const int jstart = (d->gamma && i == 0) ? 0 : -nj;
        for (int j = jstart; j <= nj; ++j) {
            const int kstart = (d->gamma && i == 0 && j == 0) ? 0 : -nk;
            for (int k = kstart; k <= nk; ++k) {
                const double tt = gnorm2(d, i, j, k);
                if (tt <= d->gcutms) {
                    mu.push_back(i);
                    mu.push_back(j);
                    mu.push_back(k);
                    g2.push_back(tt > EPS8 ? tt : 0.0);
                }
            }
        }
    }
    const int n = (int)g2.size();
    std::vector<int> igsrt(std::max(n, 1), 0);
    heap_sort_eps(n, g2.data(), igsrt.data(), EPS8);
    d->mill.resize((size_t)n * 3);
    d->g.resize((size_t)n * 3);
    d->gg.resize(n);
    int ngm = n;
    for (int ng = 0; ng < n; ++ng) {
        const int o = igsrt[ng] - 1;
        const int mi[3] = {mu[3 * o], mu[3 * o + 1], mu[3 * o + 2]};
        double s = 0.0;
        for (int c = 0; c < 3; ++c) {
            d->mill[3 * (size_t)ng + c] = mi[c];
            const double gc = (mi[0] * d->bg[c][0] + mi[1] * d->bg[c][1]) + mi[2] * d->bg[c][2];
            d->g[3 * (size_t)ng + c] = gc;
            s += gc * gc;
        }
        d->gg[ng] = s;
        if (s > d->gcutms && ngm == n) ngm = ng;  // keep the prefix with gg <= gcutms
    }
    d->ngm = ngm;
    d->mill.resize((size_t)ngm * 3);
    d->g.resize((size_t)ngm * 3);
    d->gg.resize(ngm);
    d->gstart = (ngm > 0 && d->gg[0] <= EPS8) ? 2 : 1;
    d->nl.resize(ngm);
    d->nlm.clear();
    if (d->gamma) d->nlm.resize(ngm);
    for (int ng = 0; ng < ngm; ++ng) {
        const int* m = &d->mill[3 * (size_t)ng];
        d->nl[ng] = 1 + wrap(m[0], d->nr1) + wrap(m[1], d->nr2) * d->nr1 + wrap(m[2], d->nr3) * d->nr1 * d->nr2;
        if (d->gamma)
            d->nlm[ng] = 1 + wrap(-m[0], d->nr1) + wrap(-m[1], d->nr2) * d->nr1 + wrap(-m[2], d->nr3) * d->nr1 * d->nr2;
    }
}

// gk_sort + g2_kin over a list of G-vectors (global indices `gi` into d->g; gi == nullptr: all ngm)
void gk_sort(const qeb_desc* d, const int* gi, int ng, std::vector<int>& igk, std::vector<double>& g2kin)
{
    std::vector<double> gk;
    igk.clear();
    for (int i = 0; i < ng; ++i) {
        const double* g = &d->g[3 * (size_t)(gi ? gi[i] : i)];
        const double kx = g[0] + d->xk[0], ky = g[1] + d->xk[1], kz = g[2] + d->xk[2];
        double q = (kx * kx + ky * ky) + kz * kz;
        if (q <= EPS8) q = 0.0;
        if (q <= d->gcutw) {
            gk.push_back(q);
            igk.push_back(i + 1);
        }
    }
    const double k2 = (d->xk[0] * d->xk[0] + d->xk[1] * d->xk[1]) + d->xk[2] * d->xk[2];
    if (k2 > EPS8 && !igk.empty()) heap_sort_eps((int)igk.size(), gk.data(), igk.data(), EPS8);
    g2kin.resize(igk.size());
    for (size_t i = 0; i < igk.size(); ++i) {
        const double* g = &d->g[3 * (size_t)(gi ? gi[igk[i] - 1] : igk[i] - 1)];
        const double kx = g[0] + d->xk[0], ky = g[1] + d->xk[1], kz = g[2] + d->xk[2];
        g2kin[i] = ((kx * kx + ky * ky) + kz * kz) * d->tpiba2;
    }
}

// serial isind / iplw of the wave sticks (fft_types.f90:814-851) with the Gamma mirror (stick_base.f90:594)
void wave_sticks(qeb_desc* d)
{
    int ub[3];
    std::vector<int> st;
    column_counts(d, d->gkcut, ub, st);
    const int n2 = 2 * ub[1] + 1;
    d->isind.assign((size_t)d->nr1 * d->nr2, 0);
    d->iplw.assign(d->nr1, 0);
    for (int i1 = -ub[0]; i1 <= ub[0]; ++i1)
        for (int i2 = -ub[1]; i2 <= ub[1]; ++i2) {
            bool has = st[(size_t)(i1 + ub[0]) * n2 + (i2 + ub[1])] > 0;
            if (d->gamma && !has) has = st[(size_t)(-i1 + ub[0]) * n2 + (-i2 + ub[1])] > 0;
            if (!has) continue;
            const int m1 = wrap(i1, d->nr1), m2 = wrap(i2, d->nr2);
            d->isind[m1 + (size_t)m2 * d->nr1] = 1;
            d->iplw[m1] = 1;
        }
    d->nst = 0;
    for (int v : d->isind) d->nst += v;
}

}  // namespace

namespace {

// The persistent sticks map of the reference (stick_base.f90:28-53): column numbering, processing order and owners
// survive from the wave-stick call of get_sticks to the smooth-stick call.
struct SticksMap {
    std::map<std::pair<int, int>, int> index;   // (i1,i2) -> stick number (1-based)
    std::vector<std::pair<int, int>> cols;
    std::vector<int> idx;                       // processing order
    std::map<std::pair<int, int>, int> stown;   // owner (1-based)
};

// get_sticks (stick_base.f90:549-601) for the slab decomposition (nyfft = 1): returns the owner map of the columns
// relevant at this cutoff, sticks per rank (ncp) and G-vectors per rank (ngp)
void get_sticks(const qeb_desc* d, SticksMap& sm, double gcut, int nproc, std::map<std::pair<int, int>, int>& owner,
                std::vector<int>& ncp, std::vector<long>& ngp)
{
    int ub[3];
    std::vector<int> st;
    column_counts(d, gcut, ub, st);
    const int n1 = 2 * ub[0] + 1, n2 = 2 * ub[1] + 1;
    auto cnt = [&](int i1, int i2) { return st[(size_t)(i1 + ub[0]) * n2 + (i2 + ub[1])]; };
    // sticks_map_index: j2 outer, j1 inner; non-negative indices first, then the negative ones
    for (int j2 = 0; j2 < n2; ++j2) {
        const int i2 = j2 <= ub[1] ? j2 : -ub[1] + (j2 - ub[1]) - 1;
        for (int j1 = 0; j1 < n1; ++j1) {
            const int i1 = j1 <= ub[0] ? j1 : -ub[0] + (j1 - ub[0]) - 1;
            if (cnt(i1, i2) > 0 && !sm.index.count({i1, i2})) {
                sm.cols.push_back({i1, i2});
                sm.index[{i1, i2}] = (int)sm.cols.size();
            }
        }
    }
    const int nct = (int)sm.cols.size();
    std::vector<int> ngc(nct);
    for (int mc = 0; mc < nct; ++mc) ngc[mc] = cnt(sm.cols[mc].first, sm.cols[mc].second);
    // sticks_sort_new: sticks not ordered yet, by decreasing length (hpsort orders equal keys by index)
    {
        std::vector<char> have(nct + 1, 0);
        for (int mc : sm.idx) have[mc] = 1;
        std::vector<int> fresh;
        for (int mc = 1; mc <= nct; ++mc)
            if (!have[mc] && ngc[mc - 1] > 0) fresh.push_back(mc);
        if (nproc > 1)
            std::stable_sort(fresh.begin(), fresh.end(), [&](int a, int b) { return ngc[a - 1] > ngc[b - 1]; });
        sm.idx.insert(sm.idx.end(), fresh.begin(), fresh.end());
    }
    // sticks_dist_new: owner = rank with the fewest G-vectors, then the fewest sticks, then the lowest index
    ncp.assign(nproc, 0);
    ngp.assign(nproc, 0);
    for (int mc : sm.idx) {
        const auto c = sm.cols[mc - 1];
        if (d->gamma && (c.first < 0 || (c.first == 0 && c.second < 0))) continue;
        const int n = ngc[mc - 1];
        if (n <= 0) continue;
        auto it = sm.stown.find(c);
        if (it == sm.stown.end()) {
            int jj = 0;
            for (int j = 0; j < nproc; ++j)
                if (ngp[j] < ngp[jj] || (ngp[j] == ngp[jj] && ncp[j] < ncp[jj])) jj = j;
            it = sm.stown.insert({c, jj + 1}).first;
        }
        ncp[it->second - 1] += 1;
        ngp[it->second - 1] += n;
    }
    if (d->gamma) {  // pair every stick with its -G partner on the same rank
        for (int mc : sm.idx) {
            const int n = ngc[mc - 1];
            if (n < 1) continue;
            const auto c = sm.cols[mc - 1];
            auto it = sm.stown.find(c);
            if (it == sm.stown.end()) continue;
            if (c.first == 0 && c.second == 0) {
                ngp[it->second - 1] += n - 1;
            } else {
                sm.stown[{-c.first, -c.second}] = it->second;
                ncp[it->second - 1] += 1;
                ngp[it->second - 1] += n;
            }
        }
    }
    owner.clear();
    for (int mc : sm.idx) {
        if (ngc[mc - 1] <= 0) continue;
        const auto c = sm.cols[mc - 1];
        const int o = sm.stown[c];
        owner[c] = o;
        if (d->gamma) owner[{-c.first, -c.second}] = o;
    }
}

// fft_type_init('wave') + fft_type_set (slab) + ggen/fft_set_nl/gk_sort restricted to each rank
int distribute(qeb_desc* d, int nproc)
{
    if (nproc < 1 || d->nr3 < nproc) {
        d->err = "there are processes with no planes (fft_types.f90:633-635)";
        return 1;
    }
    SticksMap sm;
    std::map<std::pair<int, int>, int> stw, st;
    std::vector<int> ncpw, ncp;
    std::vector<long> ngpw, ngp;
    get_sticks(d, sm, d->gkcut, nproc, stw, ncpw, ngpw);
    get_sticks(d, sm, d->gcutms, nproc, st, ncp, ngp);
    const int nr1 = d->nr1, nr2 = d->nr2, nr3 = d->nr3, ncol = nr1 * nr2;
    d->nproc = nproc;
    d->nr3p.assign(nproc, nr3 / nproc);
    for (int i = 0; i < nr3 % nproc; ++i) d->nr3p[i] += 1;
    std::vector<int> own(ncol, 0);  // +owner: wave stick, -owner: other smooth stick
    for (auto& kv : st) {
        const int m1 = wrap(kv.first.first, nr1), m2 = wrap(kv.first.second, nr2);
        own[m1 + m2 * nr1] = stw.count(kv.first) ? kv.second : -kv.second;
    }
    d->nsp = ncp;
    d->iss.assign(nproc, 0);
    for (int p = 1; p < nproc; ++p) d->iss[p] = d->iss[p - 1] + ncp[p - 1];
    int ntot = 0;
    for (int v : ncp) ntot += v;
    d->ismap.assign(ntot, 0);
    std::vector<int> cnt(nproc, 0);
    std::vector<int> local((size_t)ncol, 0);  // isind of the owning rank
    for (int sign = 1; sign >= -1; sign -= 2) {
        for (int pos = 0; pos < ncol; ++pos) {
            const int ip = own[pos] * sign;
            if (ip > 0) {
                cnt[ip - 1] += 1;
                d->ismap[cnt[ip - 1] - 1 + d->iss[ip - 1]] = pos + 1;
                local[pos] = cnt[ip - 1];
            }
        }
        if (sign == 1) d->nsw = cnt;
    }
    for (int p = 0; p < nproc; ++p)
        if (d->nsw[p] != ncpw[p] || cnt[p] != ncp[p]) {
            d->err = "inconsistent number of sticks (fft_types.f90:757-762)";
            return 2;
        }
    // per-rank G lists in the global order, stick-major nl/nlm, igk, g2kin
    std::vector<int> row_of_g(d->ngm, -1);
    for (int r = 0; r < d->npw; ++r) row_of_g[d->igk[r] - 1] = r;
    d->ranks.assign(nproc, RankData());
    for (int ng = 0; ng < d->ngm; ++ng) {
        const int* m = &d->mill[3 * (size_t)ng];
        const int pos = wrap(m[0], nr1) + wrap(m[1], nr2) * nr1;
        const int o = abs(own[pos]);
        if (o < 1) {
            d->err = "G-vector on an unowned stick";
            return 3;
        }
        RankData& R = d->ranks[o - 1];
        R.ig_l2g.push_back(ng);
        R.nl.push_back(1 + wrap(m[2], nr3) + (local[pos] - 1) * nr3);
        if (d->gamma) {
            const int posm = wrap(-m[0], nr1) + wrap(-m[1], nr2) * nr1;
            if (abs(own[posm]) != o) {
                d->err = "Gamma: -G stick on another rank";
                return 4;
            }
            R.nlm.push_back(1 + wrap(-m[2], nr3) + (local[posm] - 1) * nr3);
        }
    }
    for (int p = 0; p < nproc; ++p) {
        RankData& R = d->ranks[p];
        gk_sort(d, R.ig_l2g.data(), (int)R.ig_l2g.size(), R.igk, R.g2kin);
        R.pw_l2g.resize(R.igk.size());
        for (size_t i = 0; i < R.igk.size(); ++i) R.pw_l2g[i] = row_of_g[R.ig_l2g[R.igk[i] - 1]];
        R.gstart = (!R.ig_l2g.empty() && d->gg[R.ig_l2g[0]] <= EPS8) ? 2 : 1;
    }
    return 0;
}

}  // namespace

namespace {
template <class T>
int copy_out(const std::vector<T>& v, void* out)
{
    if (!v.empty()) memcpy(out, v.data(), v.size() * sizeof(T));
    return 0;
}
}  // namespace

extern "C" {

int qeb_good_fft_order(int nr) { return nr < 1 ? -1 : good_order(nr); }

int qeb_desc_create(qeb_desc** out, const double* at_cols, double alat, double ecutwfc, double ecutrho,
                    const double* xk, int gamma_only)
{
    if (!out || !at_cols || alat <= 0.0 || ecutwfc <= 0.0) return 1;
    *out = nullptr;
    qeb_desc* d = new qeb_desc();
    for (int c = 0; c < 3; ++c)
        for (int v = 0; v < 3; ++v) d->at[c][v] = at_cols[3 * c + v];  // row-major 3x3 whose COLUMNS are a_v
    recips(d);
    d->alat = alat;
    const double tpiba = 2.0 * M_PI / alat;
    d->tpiba2 = tpiba * tpiba;
    d->gcutw = ecutwfc / d->tpiba2;
    d->gcutms = 4.0 * d->gcutw;
    if (ecutrho > 0.0 && ecutrho < 4.0 * ecutwfc) d->gcutms = ecutrho / d->tpiba2;
    for (int c = 0; c < 3; ++c) d->xk[c] = xk ? xk[c] : 0.0;
    const double knorm = sqrt((d->xk[0] * d->xk[0] + d->xk[1] * d->xk[1]) + d->xk[2] * d->xk[2]);
    d->gamma = gamma_only != 0;
    if (d->gamma && knorm > 0.0) {
        delete d;
        return 2;  // gamma_only requires k = 0
    }
    const double sq = sqrt(d->gcutw) + knorm;
    d->gkcut = sq * sq;  // PW/src/data_structure.f90:59-64
    if (grid_dims(d, d->gcutms)) {
        delete d;
        return 3;
    }
    gen_gvectors(d);
    gk_sort(d, nullptr, d->ngm, d->igk, d->g2kin);
    d->npw = (int)d->igk.size();
    wave_sticks(d);
    *out = d;
    return 0;
}

int qeb_desc_destroy(qeb_desc* d)
{
    delete d;
    return 0;
}

const char* qeb_desc_error(const qeb_desc* d) { return d ? d->err.c_str() : "null descriptor"; }

int qeb_desc_dims(const qeb_desc* d, int* nr1, int* nr2, int* nr3, int* ngm, int* npw, int* nst, int* gstart)
{
    if (!d) return 1;
    if (nr1) *nr1 = d->nr1;
    if (nr2) *nr2 = d->nr2;
    if (nr3) *nr3 = d->nr3;
    if (ngm) *ngm = d->ngm;
    if (npw) *npw = d->npw;
    if (nst) *nst = d->nst;
    if (gstart) *gstart = d->gstart;
    return 0;
}

int qeb_desc_get(const qeb_desc* d, int what, int rank, void* out)
{
    if (!d || !out) return 1;
    if (rank < 0) {
        switch (what) {
        case QEB_DESC_NL: return copy_out(d->nl, out);
        case QEB_DESC_NLM: return d->gamma ? copy_out(d->nlm, out) : 2;
        case QEB_DESC_IGK: return copy_out(d->igk, out);
        case QEB_DESC_G2KIN: return copy_out(d->g2kin, out);
        case QEB_DESC_MILL: return copy_out(d->mill, out);
        case QEB_DESC_GG: return copy_out(d->gg, out);
        case QEB_DESC_ISIND: return copy_out(d->isind, out);
        case QEB_DESC_IPLW: return copy_out(d->iplw, out);
        case QEB_DESC_NR3P: return d->nproc ? copy_out(d->nr3p, out) : 3;
        case QEB_DESC_NSW: return d->nproc ? copy_out(d->nsw, out) : 3;
        case QEB_DESC_NSP: return d->nproc ? copy_out(d->nsp, out) : 3;
        case QEB_DESC_ISS: return d->nproc ? copy_out(d->iss, out) : 3;
        case QEB_DESC_ISMAP: return d->nproc ? copy_out(d->ismap, out) : 3;
        default: return 2;
        }
    }
    if (rank >= d->nproc) return 3;
    const RankData& R = d->ranks[rank];
    switch (what) {
    case QEB_DESC_NL: return copy_out(R.nl, out);
    case QEB_DESC_NLM: return d->gamma ? copy_out(R.nlm, out) : 2;
    case QEB_DESC_IGK: return copy_out(R.igk, out);
    case QEB_DESC_G2KIN: return copy_out(R.g2kin, out);
    case QEB_DESC_IG_L2G: return copy_out(R.ig_l2g, out);
    case QEB_DESC_PW_L2G: return copy_out(R.pw_l2g, out);
    default: return 2;
    }
}

int qeb_desc_distribute(qeb_desc* d, int nproc)
{
    if (!d) return 1;
    return distribute(d, nproc);
}

int qeb_desc_rank_dims(const qeb_desc* d, int rank, int* ngm, int* npw, int* gstart, int* nismap)
{
    if (!d || rank < 0 || rank >= d->nproc) return 1;
    const RankData& R = d->ranks[rank];
    if (ngm) *ngm = (int)R.ig_l2g.size();
    if (npw) *npw = (int)R.igk.size();
    if (gstart) *gstart = R.gstart;
    if (nismap) *nismap = (int)d->ismap.size();
    return 0;
}

}  // extern "C"
