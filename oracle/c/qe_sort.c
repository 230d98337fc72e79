/*
 * ORACLE (test infrastructure, not product code).
 *
 * CPU restatement of the reference's eps-aware heap sort, used by the
 * descriptor builder to reproduce QE's G-vector and k+G ordering.
 * Follows /root/reference/Modules/sort.f90:9-134 (hpsort_eps): ascending
 * heapsort of ra(1:n) carrying an index array; two keys closer than eps are
 * "equal" and are then ordered by their carried index.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
 * load this library.
 */
#include <math.h>

/* ra, ind are 0-based C arrays holding the Fortran 1-based contents. If
 * ind[0]==0 on entry the indices are initialised to 1..n (Fortran rule). */
void qeo_hpsort_eps(int n, double *ra, int *ind, double eps)
{
    int i, ir, j, l, iind;
    double rra;
    if (ind[0] == 0)
        for (i = 1; i <= n; ++i) ind[i - 1] = i;
    if (n < 2) return;
    l = n / 2 + 1;
    ir = n;
#define RA(k) ra[(k) - 1]
#define IND(k) ind[(k) - 1]
    for (;;) {
        if (l > 1) {
            l = l - 1;
            rra = RA(l);
            iind = IND(l);
        } else {
            rra = RA(ir);
            iind = IND(ir);
            RA(ir) = RA(1);
            IND(ir) = IND(1);
            ir = ir - 1;
            if (ir == 1) {
                RA(1) = rra;
                IND(1) = iind;
                break;
            }
        }
        i = l;
        j = l + l;
        while (j <= ir) {
            if (j < ir) {
                if (fabs(RA(j) - RA(j + 1)) >= eps) {
                    if (RA(j) < RA(j + 1)) j = j + 1;
                } else {
                    if (IND(j) < IND(j + 1)) j = j + 1;
                }
            }
            if (fabs(rra - RA(j)) >= eps) {
                if (rra < RA(j)) {
                    RA(i) = RA(j);
                    IND(i) = IND(j);
                    i = j;
                    j = j + j;
                } else {
                    j = ir + 1;
                }
            } else {
                if (iind < IND(j)) {
                    RA(i) = RA(j);
                    IND(i) = IND(j);
                    i = j;
                    j = j + j;
                } else {
                    j = ir + 1;
                }
            }
        }
        RA(i) = rra;
        IND(i) = iind;
    }
#undef RA
#undef IND
}
