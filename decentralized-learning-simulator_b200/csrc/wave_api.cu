// C-ABI entry points for the train / grad / test waves (see include/dls_b200.h).
#include "gnlenet.cuh"

namespace dls {
namespace fem {
int64_t train_workspace_bytes(int n_tasks, int batch, bool own_G);
int step(const gnl::Pass &p, void *workspace, float *G, int64_t ldg, float *loss_out, float **G_used,
         int64_t *ldg_used, int precision, const gnl::SgdFuse *sgd, cudaStream_t st);
int64_t eval_workspace_bytes(int n_models, int n_test);
int eval(const float *X, int64_t ldx, const int32_t *model_rows, int n_models, const uint8_t *test_images,
         const int64_t *test_labels, int n_test, int32_t *out_correct, double *out_loss_sum, float *out_logits,
         int precision, void *workspace, cudaStream_t st);
}  // namespace fem

// Fused optimizer update over the wave's rows — torch.optim.SGD semantics
// (reference: dasklearn/optimizer/sgd.py:4-14; state carry-over dasklearn/model_trainer.py:89-92):
//   g += wd * p;  buf = g (node's first ever step) else buf = mu * buf + g;  p -= lr * buf
// grid = (column blocks, n_tasks).
__global__ void __launch_bounds__(256)
sgd_update_kernel(float *__restrict__ X, int64_t ldx, float *__restrict__ M, int64_t ldm,
                  const float *__restrict__ G, int64_t ldg, const int32_t *__restrict__ task_row,
                  const int32_t *__restrict__ task_peer, const uint8_t *__restrict__ first,
                  const int32_t *__restrict__ gcount, int64_t P, float lr, float mu, float wd, int skip0, int skip1)
{
    // [skip0, skip1): parameters already updated by the weight-gradient epilogues (SgdFuse)
    const int g = blockIdx.y;
    if (gcount[g] <= 0) return;
    float *x = X + (int64_t)task_row[g] * ldx;
    const float *gr = G + (int64_t)g * ldg;
    float *mb = (M && mu != 0.f) ? M + (int64_t)task_peer[g] * ldm : nullptr;
    const bool fst = first ? (first[g] != 0) : false;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P;
         i += (int64_t)gridDim.x * blockDim.x) {
        if (i >= skip0 && i < skip1) continue;
        float gv = gr[i];
        const float pv = x[i];
        if (wd != 0.f) gv = fmaf(wd, pv, gv);
        if (mb) {
            const float b = fst ? gv : __fadd_rn(__fmul_rn(mb[i], mu), gv);
            mb[i] = b;
            gv = b;
        }
        x[i] = fmaf(-lr, gv, pv);
    }
}

__global__ void eval_desc_kernel(int32_t *gmodel, int32_t *img_idx, int32_t *gcount, int g0, int n_groups,
                                 int gs, int gpm, int n_test)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_groups * gs) return;
    const int gl = i / gs, j = i - gl * gs, g = g0 + gl;
    const int m = g / gpm, chunk = g - m * gpm;
    const int sample = chunk * gs + j;
    img_idx[i] = sample < n_test ? sample : -1;
    if (j == 0) {
        gmodel[gl] = m;
        const int rem = n_test - chunk * gs;
        gcount[gl] = rem < gs ? (rem > 0 ? rem : 0) : gs;
    }
}

// Sum per-group eval results of each model in group order (deterministic, fp64).
__global__ void eval_reduce_kernel(const float *glog_all, int gpm, int n_models, int32_t *out_correct,
                                   double *out_loss)
{
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n_models) return;
    double ls = 0.0;
    int hs = 0;
    for (int k = 0; k < gpm; ++k) {
        ls += (double)glog_all[((int64_t)m * gpm + k) * 2];
        hs += (int)glog_all[((int64_t)m * gpm + k) * 2 + 1];
    }
    out_loss[m] = ls;
    out_correct[m] = hs;
}

constexpr int EVAL_GS = 32;
constexpr int EVAL_CHUNK_GROUPS = 512;   // 16384 sample slots per pass

}  // namespace dls

using namespace dls;

static int check_rows(const void *X, int64_t ld, int64_t P)
{
    if (!X) return DLS_E_ARG;
    if (!aligned16(X) || ld % 4) return DLS_E_ALIGN;
    if (ld < P) return DLS_E_SHAPE;
    return 0;
}

// A train wave with more tasks than the GPU has SMs runs as TWO independent lanes on two streams: the lanes
// share nothing (disjoint tasks, workspaces, rows), so kernels of different kinds overlap — a tensor-bound
// convolution of one half next to a CUDA-core GroupNorm kernel of the other — and an SM never idles while
// the second wave of a one-CTA-per-SM kernel drains.
// The first lane takes a whole number of SM-fulls of tasks (148 of 256 on a B200): its one-CTA-per-task kernels
// then run in full waves, and the second lane's CTAs fill whatever the first leaves idle.  Measured at 256
// tasks: 11.13 ms per round unsplit, 11.07 ms split 128/128, 10.77 ms split 148/108.
static int split_point(int n_tasks)
{
    int sms = num_sms();                                      // of the current device
    if (sms <= 0) sms = 148;
    if (n_tasks <= sms) return n_tasks;                       // a single wave: nothing to gain
    int k = (n_tasks + sms) / (2 * sms);                      // round(n / (2 sms))
    if (k < 1) k = 1;
    const int na = k * sms;
    return na < n_tasks ? na : n_tasks;
}

// Lanes of a train wave: two by default (split_point); dls_tune(TUNE_LANES, k) forces k equal lanes (experiments).
static int lane_split(int n_tasks, int *lane_n, int *lane_off)
{
    const int k0 = tune_value(TUNE_LANES);
    if (k0 >= 1 && k0 <= gnl::MAX_LANES) {
        int k = k0;
        if (k > n_tasks) k = n_tasks > 0 ? n_tasks : 1;
        int off = 0;
        for (int h = 0; h < k; ++h) {
            lane_n[h] = n_tasks / k + (h < n_tasks % k ? 1 : 0);
            lane_off[h] = off;
            off += lane_n[h];
        }
        return k;
    }
    const int na = split_point(n_tasks);
    lane_n[0] = na; lane_off[0] = 0;
    if (na == n_tasks) return 1;
    lane_n[1] = n_tasks - na; lane_off[1] = na;
    return 2;
}

extern "C" int64_t dls_train_workspace_bytes(int32_t arch, int32_t n_tasks, int32_t batch)
{
    if (n_tasks < 0 || batch <= 0 || batch > DLS_MAX_GROUP) return DLS_E_SHAPE;
    // sized for either precision (the tensor-core path adds bf16 weight copies)
    if (arch == DLS_ARCH_GNLENET) {
        int lane_n[gnl::MAX_LANES], lane_off[gnl::MAX_LANES];
        const int n_lanes = lane_split(n_tasks, lane_n, lane_off);
        int64_t total = 0;
        for (int h = 0; h < n_lanes; ++h) total += gnl::workspace_bytes(lane_n[h], batch, lane_n[h], true, true, 1);
        const int64_t whole = gnl::workspace_bytes(n_tasks, batch, n_tasks, true, true, 1);
        return whole > total ? whole : total;
    }
    if (arch == DLS_ARCH_FEMNIST) return fem::train_workspace_bytes(n_tasks, batch, true);
    return DLS_E_SHAPE;
}

static int gnl_step(const gnl::Pass &p, gnl::Workspace &ws, float *loss_out, cudaStream_t st, int lane = 0)
{
    int rc;
    if ((rc = gnl::prep_weights(p, ws, true, st))) return rc;
    if ((rc = gnl::forward(p, ws, true, st))) return rc;
    if ((rc = gnl::classifier(p, ws, 0, loss_out, nullptr, 0, st))) return rc;
    if ((rc = gnl::backward(p, ws, st, lane))) return rc;
    return 0;
}



extern "C" int dls_train_wave(int32_t arch, float *X, int64_t ldx, float *M, int64_t ldm,
                              const int32_t *task_row, const int32_t *task_peer,
                              const uint8_t *mom_first, const uint8_t *images, const int64_t *labels,
                              const int32_t *mb_index, const int32_t *mb_count, int32_t n_tasks,
                              int32_t n_steps, int32_t batch, float lr, float mu, float wd,
                              float *loss_out, int32_t precision, void *workspace,
                              int64_t workspace_bytes, void *stream)
{
    if (arch != DLS_ARCH_GNLENET && arch != DLS_ARCH_FEMNIST) return DLS_E_SHAPE;
    if (precision != 0 && precision != 1) return DLS_E_SHAPE;
    if (arch == DLS_ARCH_FEMNIST) {
        // FEMNIST CNN: precision 1 runs conv2 on tensor cores (femnist.cu)
        if (!task_row || !task_peer || !images || !labels || !mb_index || !mb_count || !workspace || n_tasks < 0 ||
            n_steps < 0)
            return DLS_E_ARG;
        if (batch <= 0 || batch > DLS_MAX_GROUP) return DLS_E_SHAPE;
        int rcf = check_rows(X, ldx, DLS_FEMNIST_P);
        if (rcf) return rcf;
        if (mu != 0.f) {
            if ((rcf = check_rows(M, ldm, DLS_FEMNIST_P))) return rcf;
            if (!mom_first) return DLS_E_ARG;
        }
        if (n_tasks == 0 || n_steps == 0) return 0;
        if (workspace_bytes < fem::train_workspace_bytes(n_tasks, batch, true)) return DLS_E_WS;
        cudaStream_t stf = (cudaStream_t)stream;
        for (int s = 0; s < n_steps; ++s) {
            gnl::Pass p;
            p.n_groups = n_tasks; p.gs = batch; p.n_models = n_tasks; p.gmodel = nullptr; p.mrow = task_row;
            p.img_idx = mb_index + (int64_t)s * n_tasks * batch; p.gcount = mb_count + (int64_t)s * n_tasks;
            p.images = images; p.labels = labels; p.X = X; p.ldx = ldx;
            float *G = nullptr;
            int64_t ldg = 0;
            // precision 1: the optimizer step of fc1.weight (95 % of P) happens in its weight-gradient epilogue
            const bool fuse = precision == 1 && !tune_value(TUNE_NO_SGD_FUSE);
            gnl::SgdFuse o;
            if (fuse) {
                o.X = X; o.ldx = ldx; o.M = M; o.ldm = ldm; o.task_row = task_row; o.task_peer = task_peer;
                o.first = s == 0 ? mom_first : nullptr; o.lr = lr; o.mu = mu; o.wd = wd;
            }
            if ((rcf = fem::step(p, workspace, nullptr, 0, loss_out ? loss_out + (int64_t)s * n_tasks : nullptr, &G,
                                 &ldg, precision, fuse ? &o : nullptr, stf)))
                return rcf;
            if (fuse) {
                constexpr int F0 = 52096, F1 = 1657728;        // fc1.weight = [F0, F1) of the flat parameter row
                sgd_update_kernel<<<dim3(8, n_tasks), 256, 0, stf>>>(X, ldx, M, ldm, G, ldg, task_row, task_peer,
                                                                     s == 0 ? mom_first : nullptr, p.gcount, F0, lr, mu, wd, 0, 0);
                DLS_LAUNCH_CHECK();
                sgd_update_kernel<<<dim3(8, n_tasks), 256, 0, stf>>>(X + F1, ldx, M ? M + F1 : nullptr, ldm, G + F1, ldg, task_row,
                                                                     task_peer, s == 0 ? mom_first : nullptr, p.gcount,
                                                                     DLS_FEMNIST_P - F1, lr, mu, wd, 0, 0);
                DLS_LAUNCH_CHECK();
            } else {
                dim3 grid(128, n_tasks);
                sgd_update_kernel<<<grid, 256, 0, stf>>>(X, ldx, M, ldm, G, ldg, task_row, task_peer,
                                                         s == 0 ? mom_first : nullptr, p.gcount, DLS_FEMNIST_P, lr, mu, wd, 0, 0);
                DLS_LAUNCH_CHECK();
            }
        }
        return 0;
    }
    if (!task_row || !task_peer || !images || !labels || !mb_index || !mb_count || !workspace ||
        n_tasks < 0 || n_steps < 0)
        return DLS_E_ARG;
    if (batch <= 0 || batch > DLS_MAX_GROUP) return DLS_E_SHAPE;
    int rc = check_rows(X, ldx, gnl::P);
    if (rc) return rc;
    if (mu != 0.f) {
        if ((rc = check_rows(M, ldm, gnl::P))) return rc;
        if (!mom_first) return DLS_E_ARG;
    }
    if (n_tasks == 0 || n_steps == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    int lane_n[gnl::MAX_LANES], lane_off[gnl::MAX_LANES];
    const int n_lanes = lane_split(n_tasks, lane_n, lane_off);
    int64_t lane_bytes[gnl::MAX_LANES], total = 0;
    for (int h = 0; h < n_lanes; ++h) {
        lane_bytes[h] = gnl::workspace_bytes(lane_n[h], batch, lane_n[h], true, true, precision);
        total += lane_bytes[h];
    }
    if (workspace_bytes < total) return DLS_E_WS;
    gnl::Workspace ws[gnl::MAX_LANES];
    cudaStream_t lane_stream[gnl::MAX_LANES];
    gnl::SideStream *lane_ss[gnl::MAX_LANES];
    char *wbase = reinterpret_cast<char *>(workspace);
    for (int h = 0; h < n_lanes; ++h) {
        gnl::carve(ws[h], wbase, lane_n[h], batch, lane_n[h], true, true, precision);
        wbase += lane_bytes[h];
        lane_stream[h] = st;
        lane_ss[h] = nullptr;
        if (h > 0) {                                                  // lanes >= 1 start after everything queued on `stream`
            if ((rc = gnl::side_stream(&lane_ss[h], 2 * h + 1))) return rc;
            lane_stream[h] = lane_ss[h]->s;
            cudaError_t e = cudaEventRecord(lane_ss[h]->fork, st);
            if (e == cudaSuccess) e = cudaStreamWaitEvent(lane_ss[h]->s, lane_ss[h]->fork, 0);
            if (e != cudaSuccess) return (int)e;
        }
    }
    for (int s = 0; s < n_steps; ++s) {
        for (int h = 0; h < n_lanes; ++h) {
            const int off = lane_off[h], nt = lane_n[h];
            cudaStream_t sh = lane_stream[h];
            gnl::Pass p;
            p.n_groups = nt; p.gs = batch; p.n_models = nt;
            p.gmodel = nullptr; p.mrow = task_row + off;
            p.img_idx = mb_index + ((int64_t)s * n_tasks + off) * batch;
            p.gcount = mb_count + (int64_t)s * n_tasks + off;
            p.images = images; p.labels = labels; p.X = X; p.ldx = ldx;
            // tensor-core path: the conv2 / conv3 weights (OFF_W2 .. OFF_B2 and OFF_W3 .. OFF_B3) are updated inside
            // their weight-gradient kernels; the update kernel below then only walks the rest of the row
            const bool fuse = precision == 1 && !tune_value(TUNE_CONV_UNGROUPED) && !tune_value(TUNE_NO_SGD_FUSE);
            ws[h].sgd = gnl::SgdFuse();
            if (fuse) {
                gnl::SgdFuse &o = ws[h].sgd;
                o.X = X; o.ldx = ldx; o.M = M; o.ldm = ldm; o.task_row = task_row + off; o.task_peer = task_peer + off;
                o.first = (s == 0 && mom_first) ? mom_first + off : nullptr;
                o.lr = lr; o.mu = mu; o.wd = wd;
            }
            if ((rc = gnl_step(p, ws[h], loss_out ? loss_out + (int64_t)s * n_tasks + off : nullptr, sh, h))) return rc;
            // two launches around the fused ranges: [0, OFF_W2) + [OFF_B2, OFF_W3) in one grid, [OFF_B3, P) in the other
            if (fuse) {
                dim3 g1(4, nt);
                sgd_update_kernel<<<g1, 256, 0, sh>>>(X, ldx, M, ldm, ws[h].G, ws[h].ldg, task_row + off, task_peer + off,
                                                      (s == 0 && mom_first) ? mom_first + off : nullptr, p.gcount,
                                                      gnl::OFF_W3, lr, mu, wd, gnl::OFF_W2, gnl::OFF_B2);
                DLS_LAUNCH_CHECK();
                sgd_update_kernel<<<g1, 256, 0, sh>>>(X + gnl::OFF_B3, ldx, M ? M + gnl::OFF_B3 : nullptr, ldm,
                                                      ws[h].G + gnl::OFF_B3, ws[h].ldg, task_row + off, task_peer + off,
                                                      (s == 0 && mom_first) ? mom_first + off : nullptr, p.gcount,
                                                      gnl::P - gnl::OFF_B3, lr, mu, wd, 0, 0);
                DLS_LAUNCH_CHECK();
                continue;
            }
            dim3 grid(32, nt);
            sgd_update_kernel<<<grid, 256, 0, sh>>>(X, ldx, M, ldm, ws[h].G, ws[h].ldg, task_row + off, task_peer + off,
                                                    (s == 0 && mom_first) ? mom_first + off : nullptr, p.gcount, gnl::P,
                                                    lr, mu, wd, 0, 0);
            DLS_LAUNCH_CHECK();
        }
    }
    for (int h = 1; h < n_lanes; ++h) {                                // `stream` continues after every lane
        cudaError_t e = cudaEventRecord(lane_ss[h]->join, lane_ss[h]->s);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(st, lane_ss[h]->join, 0);
        if (e != cudaSuccess) return (int)e;
    }
    return 0;
}

extern "C" int dls_grad_wave(int32_t arch, const float *X, int64_t ldx, float *G, int64_t ldg,
                             const int32_t *task_row, const uint8_t *images, const int64_t *labels,
                             const int32_t *mb_index, const int32_t *mb_count, int32_t n_tasks,
                             int32_t batch, float *loss_out, int32_t precision, void *workspace,
                             int64_t workspace_bytes, void *stream)
{
    if (arch != DLS_ARCH_GNLENET && arch != DLS_ARCH_FEMNIST) return DLS_E_SHAPE;
    if (precision != 0 && precision != 1) return DLS_E_SHAPE;
    if (arch == DLS_ARCH_FEMNIST) {
        if (!task_row || !images || !labels || !mb_index || !mb_count || !workspace || n_tasks < 0) return DLS_E_ARG;
        if (batch <= 0 || batch > DLS_MAX_GROUP) return DLS_E_SHAPE;
        int rcf = check_rows(X, ldx, DLS_FEMNIST_P);
        if (rcf) return rcf;
        if ((rcf = check_rows(G, ldg, DLS_FEMNIST_P))) return rcf;
        if (n_tasks == 0) return 0;
        if (workspace_bytes < fem::train_workspace_bytes(n_tasks, batch, false)) return DLS_E_WS;
        gnl::Pass p;
        p.n_groups = n_tasks; p.gs = batch; p.n_models = n_tasks; p.gmodel = nullptr; p.mrow = task_row;
        p.img_idx = mb_index; p.gcount = mb_count; p.images = images; p.labels = labels; p.X = X; p.ldx = ldx;
        return fem::step(p, workspace, G, ldg, loss_out, nullptr, nullptr, precision, nullptr, (cudaStream_t)stream);
    }
    if (!task_row || !images || !labels || !mb_index || !mb_count || !workspace || n_tasks < 0)
        return DLS_E_ARG;
    if (batch <= 0 || batch > DLS_MAX_GROUP) return DLS_E_SHAPE;
    int rc = check_rows(X, ldx, gnl::P);
    if (rc) return rc;
    if ((rc = check_rows(G, ldg, gnl::P))) return rc;
    if (n_tasks == 0) return 0;
    if (workspace_bytes < gnl::workspace_bytes(n_tasks, batch, n_tasks, true, false, precision)) return DLS_E_WS;
    gnl::Workspace ws;
    gnl::carve(ws, workspace, n_tasks, batch, n_tasks, true, false, precision);
    ws.G = G; ws.ldg = ldg;
    gnl::Pass p;
    p.n_groups = n_tasks; p.gs = batch; p.n_models = n_tasks;
    p.gmodel = nullptr; p.mrow = task_row; p.img_idx = mb_index; p.gcount = mb_count;
    p.images = images; p.labels = labels; p.X = X; p.ldx = ldx;
    return gnl_step(p, ws, loss_out, (cudaStream_t)stream);
}

extern "C" int64_t dls_eval_workspace_bytes(int32_t arch, int32_t n_models, int32_t n_test)
{
    if (arch == DLS_ARCH_FEMNIST && n_models >= 0 && n_test >= 0) return fem::eval_workspace_bytes(n_models, n_test);
    if (arch != DLS_ARCH_GNLENET || n_models < 0 || n_test < 0) return DLS_E_SHAPE;
    const int gpm = (n_test + EVAL_GS - 1) / EVAL_GS;
    const int64_t total_groups = (int64_t)n_models * gpm;
    const int chunk = (int)(total_groups < EVAL_CHUNK_GROUPS ? total_groups : EVAL_CHUNK_GROUPS);
    return gnl::workspace_bytes(chunk > 0 ? chunk : 1, EVAL_GS, n_models, false, false, 1) +
           total_groups * 2 * (int64_t)sizeof(float) + 1024;
}

extern "C" int dls_eval_wave(int32_t arch, const float *X, int64_t ldx, const int32_t *model_rows,
                             int32_t n_models, const uint8_t *test_images, const int64_t *test_labels,
                             int32_t n_test, int32_t *out_correct, double *out_loss_sum,
                             float *out_logits, int32_t precision, void *workspace,
                             int64_t workspace_bytes, void *stream)
{
    if (arch != DLS_ARCH_GNLENET && arch != DLS_ARCH_FEMNIST) return DLS_E_SHAPE;
    if (precision != 0 && precision != 1) return DLS_E_SHAPE;
    if (arch == DLS_ARCH_FEMNIST) {
        if (!model_rows || !test_images || !test_labels || !out_correct || !out_loss_sum || !workspace ||
            n_models < 0 || n_test < 0)
            return DLS_E_ARG;
        int rcf = check_rows(X, ldx, DLS_FEMNIST_P);
        if (rcf) return rcf;
        if (n_models == 0 || n_test == 0) return 0;
        if (workspace_bytes < fem::eval_workspace_bytes(n_models, n_test)) return DLS_E_WS;
        return fem::eval(X, ldx, model_rows, n_models, test_images, test_labels, n_test, out_correct, out_loss_sum,
                         out_logits, precision, workspace, (cudaStream_t)stream);
    }
    if (!model_rows || !test_images || !test_labels || !out_correct || !out_loss_sum || !workspace ||
        n_models < 0 || n_test < 0)
        return DLS_E_ARG;
    int rc = check_rows(X, ldx, gnl::P);
    if (rc) return rc;
    if (n_models == 0 || n_test == 0) return 0;
    if (workspace_bytes < dls_eval_workspace_bytes(arch, n_models, n_test)) return DLS_E_WS;
    cudaStream_t st = (cudaStream_t)stream;
    const int gpm = (n_test + EVAL_GS - 1) / EVAL_GS;
    const int64_t total_groups = (int64_t)n_models * gpm;
    const int chunk = (int)(total_groups < EVAL_CHUNK_GROUPS ? total_groups : EVAL_CHUNK_GROUPS);
    gnl::Workspace ws;
    const int64_t used = gnl::carve(ws, workspace, chunk, EVAL_GS, n_models, false, false, precision);
    float *glog_all = reinterpret_cast<float *>(reinterpret_cast<char *>(workspace) + used);
    int32_t *d_img_idx = ws.desc_i32;
    int32_t *d_gmodel = ws.desc_i32 + (int64_t)chunk * EVAL_GS;
    int32_t *d_gcount = d_gmodel + chunk;
    gnl::Pass p;
    p.gs = EVAL_GS; p.n_models = n_models; p.gmodel = d_gmodel; p.mrow = model_rows;
    p.img_idx = d_img_idx; p.gcount = d_gcount; p.images = test_images; p.labels = test_labels;
    p.X = X; p.ldx = ldx; p.n_groups = chunk;
    if ((rc = gnl::prep_weights(p, ws, false, st))) return rc;
    for (int64_t g0 = 0; g0 < total_groups; g0 += chunk) {
        const int ng = (int)((total_groups - g0) < chunk ? (total_groups - g0) : chunk);
        p.n_groups = ng;
        eval_desc_kernel<<<(ng * EVAL_GS + 255) / 256, 256, 0, st>>>(d_gmodel, d_img_idx, d_gcount, (int)g0, ng,
                                                                      EVAL_GS, gpm, n_test);
        DLS_LAUNCH_CHECK();
        ws.glog = glog_all + g0 * 2;
        if ((rc = gnl::forward(p, ws, false, st))) return rc;
        if ((rc = gnl::classifier(p, ws, 1, nullptr, out_logits, n_test * gnl::NCLS, st))) return rc;
    }
    eval_reduce_kernel<<<(n_models + 127) / 128, 128, 0, st>>>(glog_all, gpm, n_models, out_correct, out_loss_sum);
    DLS_LAUNCH_CHECK();
    return 0;
}
