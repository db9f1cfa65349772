/*
 * ORACLE — test infrastructure only.  CPU restatement of the reference's parameter averaging
 * for the `aggregate` task.  Nothing in the product path may link or call this file; only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it.
 *
 * Follows FedAvg.aggregate, /root/reference/dasklearn/gradient_aggregation/fedavg.py:12-26:
 *     center = deepcopy(models[0]); p.mul_(0)            -> acc = x_0 * 0.0f   (keeps sign of zero / NaN)
 *     for m, w in zip(models, weights): c.add_(w * p)     -> tmp = fl32(w) * x_j (rounded); acc = acc + tmp (rounded)
 * i.e. list order, separate multiply and add, NO fused multiply-add.  Build with
 * -ffp-contract=off so the compiler cannot contract the two roundings into an FMA.
 *
 * Parity pin: checked bit-for-bit against the reference's own FedAvg.aggregate through the
 * fixture tests/golden/fedavg_kat.npz (made by tests/golden/make_golden.py running the
 * unmodified reference in the build container).
 */
#include <stddef.h>
#include <stdint.h>

/* One aggregate: y[0..P) = sum_j w[j] * X[cols[j]*ldx + 0..P) in list order. */
void oracle_fedavg_row(const float *X, int64_t ldx, const int32_t *cols, const float *w,
                       int fan_in, int64_t P, float *y)
{
    if (fan_in <= 0) { for (int64_t i = 0; i < P; ++i) y[i] = 0.0f; return; }
    const float *x0 = X + (int64_t)cols[0] * ldx;
    for (int64_t i = 0; i < P; ++i) y[i] = x0[i] * 0.0f;
    for (int j = 0; j < fan_in; ++j) {
        const float *xj = X + (int64_t)cols[j] * ldx;
        const float wj = w[j];
        for (int64_t i = 0; i < P; ++i) {
            volatile float tmp = wj * xj[i];   /* force the product to be rounded to fp32 */
            y[i] = y[i] + tmp;
        }
    }
}

/* A whole aggregate wave in CSR form: row r of Y (at out_rows[r], or r when out_rows == NULL). */
void oracle_fedavg_csr(const float *X, int64_t ldx, float *Y, int64_t ldy, const int32_t *rowptr,
                       const int32_t *colidx, const float *wval, const int32_t *out_rows,
                       int n_rows, int64_t P)
{
    for (int r = 0; r < n_rows; ++r) {
        int b = rowptr[r], e = rowptr[r + 1];
        int64_t orow = out_rows ? out_rows[r] : r;
        oracle_fedavg_row(X, ldx, colidx + b, wval + b, e - b, P, Y + orow * ldy);
    }
}
