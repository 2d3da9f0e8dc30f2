// PyTorch extension form of the placement ops (north star: "a thin C-ABI/pybind layer, exposed as a PyTorch
// extension taking int64 and float64 tensors").
//
//   torch.ops.k8s_place.snapshot_place(alloc_cpu_milli, alloc_mem_bytes, pod_count, max_pods, ready, cpu_usage_pct,
//                                      mem_usage_pct, req_cpu_milli, req_mem_bytes, strategy_ids?, strategy, filter_mode)
//       -> (node_idx int64[P], score float64[P])
//   torch.ops.k8s_place.snapshot_place_out(... same ..., out_idx, out_score, workspace) -> ()      caller-owned buffers
//   torch.ops.k8s_place.snapshot_place_pruned(... same inputs ..., probe_depth) -> (node_idx, score)   first-feasible probe (option)
//   torch.ops.k8s_place.greedy_place(... same inputs ...) -> (node_idx, score, pod_count, cpu_usage_pct, mem_usage_pct)
//   torch.ops.k8s_place.node_scores(7 node columns, strategy, filter_mode) -> (score, free_cpu, free_mem, eligible uint8)
//   torch.ops.k8s_place.snapshot_workspace_bytes(n_nodes, n_pods, per_pod) / greedy_workspace_bytes(n_nodes, n_pods)
//
// Nothing is computed here: every op checks its tensors, takes torch's current CUDA stream and calls the C ABI of
// libk8s_place_b200.so (include/k8s_place_b200.h), which this module links.  Registered for the CUDA dispatch key
// only -- there is no CPU implementation to fall back to.
#include <ATen/ATen.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/library.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <tuple>

#include "../../include/k8s_place_b200.h"

namespace {

using at::Tensor;

// error text is formatted here with snprintf and handed to TORCH_CHECK as one string
[[noreturn]] void fail(const char *fmt, ...)
{
    char buf[640];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    TORCH_CHECK(false, std::string(buf));
    abort();
}

const void *col(const Tensor &t, const char *name, at::ScalarType dtype, const c10::Device &dev, int64_t n)
{
    if (!t.is_cuda()) fail("%s: expected a CUDA tensor", name);
    if (t.device() != dev) fail("%s: on %s, expected %s", name, t.device().str().c_str(), dev.str().c_str());
    if (t.scalar_type() != dtype) fail("%s: dtype %s, expected %s", name, c10::toString(t.scalar_type()), c10::toString(dtype));
    if (t.dim() != 1 || t.numel() != n) fail("%s: expected shape (%lld,), got %lld element(s) in %d dimension(s)", name, (long long)n,
                                              (long long)t.numel(), (int)t.dim());
    if (!t.is_contiguous()) fail("%s: must be contiguous", name);
    return t.data_ptr();
}

struct Tables {
    kp_node_table nodes;
    kp_pod_table pods;
    c10::Device dev;
    bool per_pod;
};

kp_node_table node_table(const Tensor &ac, const Tensor &am, const Tensor &pc, const Tensor &mp, const Tensor &ready, const Tensor &cu,
                         const Tensor &mu)
{
    if (!pc.is_cuda()) fail("pod_count: expected a CUDA tensor");
    const c10::Device dev = pc.device();
    const int64_t n = pc.numel();
    kp_node_table t{};
    t.n_nodes = n;
    t.alloc_cpu_milli = static_cast<const int64_t *>(col(ac, "alloc_cpu_milli", at::kLong, dev, n));
    t.alloc_mem_bytes = static_cast<const int64_t *>(col(am, "alloc_mem_bytes", at::kLong, dev, n));
    t.pod_count = static_cast<const int64_t *>(col(pc, "pod_count", at::kLong, dev, n));
    t.max_pods = static_cast<const int64_t *>(col(mp, "max_pods", at::kLong, dev, n));
    t.ready = static_cast<const uint8_t *>(col(ready, "ready", at::kByte, dev, n));
    t.cpu_usage_pct = static_cast<const double *>(col(cu, "cpu_usage_pct", at::kDouble, dev, n));
    t.mem_usage_pct = static_cast<const double *>(col(mu, "mem_usage_pct", at::kDouble, dev, n));
    return t;
}

Tables tables(const Tensor &ac, const Tensor &am, const Tensor &pc, const Tensor &mp, const Tensor &ready, const Tensor &cu, const Tensor &mu,
              const Tensor &rc, const Tensor &rm, const c10::optional<Tensor> &sid)
{
    Tables t{node_table(ac, am, pc, mp, ready, cu, mu), kp_pod_table{}, pc.device(), sid.has_value()};
    const int64_t p = rc.numel();
    t.pods.n_pods = p;
    t.pods.req_cpu_milli = static_cast<const int64_t *>(col(rc, "req_cpu_milli", at::kLong, t.dev, p));
    t.pods.req_mem_bytes = static_cast<const int64_t *>(col(rm, "req_mem_bytes", at::kLong, t.dev, p));
    t.pods.strategy_ids = sid.has_value() ? static_cast<const int64_t *>(col(*sid, "strategy_ids", at::kLong, t.dev, p)) : nullptr;
    return t;
}

void check(int status, const char *where)
{
    if (status != KP_OK) fail("%s: status %d (%s: %s)", where, status, kp_status_string(status), kp_last_error());
}

void snapshot_place_out(const Tensor &ac, const Tensor &am, const Tensor &pc, const Tensor &mp, const Tensor &ready, const Tensor &cu,
                        const Tensor &mu, const Tensor &rc, const Tensor &rm, const c10::optional<Tensor> &sid, int64_t strategy,
                        int64_t filter_mode, Tensor out_idx, Tensor out_score, Tensor workspace)
{
    const Tables t = tables(ac, am, pc, mp, ready, cu, mu, rc, rm, sid);
    col(out_idx, "out_idx", at::kLong, t.dev, t.pods.n_pods);
    col(out_score, "out_score", at::kDouble, t.dev, t.pods.n_pods);
    if (!(workspace.is_cuda() && workspace.device() == t.dev && workspace.scalar_type() == at::kByte && workspace.is_contiguous()))
        fail("workspace: expected a contiguous uint8 tensor on %s", t.dev.str().c_str());
    const c10::cuda::CUDAGuard guard(t.dev);
    check(kp_snapshot_place(&t.nodes, &t.pods, (int)strategy, (int)filter_mode, out_idx.data_ptr<int64_t>(), out_score.data_ptr<double>(),
                            workspace.data_ptr(), (size_t)workspace.numel(), c10::cuda::getCurrentCUDAStream(t.dev.index()).stream()),
          "kp_snapshot_place");
}

std::tuple<Tensor, Tensor> snapshot_place(const Tensor &ac, const Tensor &am, const Tensor &pc, const Tensor &mp, const Tensor &ready,
                                          const Tensor &cu, const Tensor &mu, const Tensor &rc, const Tensor &rm,
                                          const c10::optional<Tensor> &sid, int64_t strategy, int64_t filter_mode)
{
    const int64_t p = rc.numel();
    Tensor out_idx = at::empty({p}, rc.options().dtype(at::kLong));
    Tensor out_score = at::empty({p}, rc.options().dtype(at::kDouble));
    const int64_t ws = (int64_t)kp_snapshot_workspace_bytes(pc.numel(), p, sid.has_value());
    Tensor workspace = at::empty({ws < 256 ? 256 : ws}, rc.options().dtype(at::kByte));
    snapshot_place_out(ac, am, pc, mp, ready, cu, mu, rc, rm, sid, strategy, filter_mode, out_idx, out_score, workspace);
    return {out_idx, out_score};
}

std::tuple<Tensor, Tensor, Tensor, Tensor, Tensor> greedy_place(const Tensor &ac, const Tensor &am, const Tensor &pc, const Tensor &mp,
                                                                const Tensor &ready, const Tensor &cu, const Tensor &mu, const Tensor &rc,
                                                                const Tensor &rm, const c10::optional<Tensor> &sid, int64_t strategy,
                                                                int64_t filter_mode, const c10::optional<Tensor> &workspace)
{
    const Tables t = tables(ac, am, pc, mp, ready, cu, mu, rc, rm, sid);
    const int64_t p = t.pods.n_pods, n = t.nodes.n_nodes;
    Tensor out_idx = at::empty({p}, rc.options().dtype(at::kLong)), out_score = at::empty({p}, rc.options().dtype(at::kDouble));
    Tensor out_pc = at::empty({n}, pc.options()), out_cu = at::empty({n}, cu.options()), out_mu = at::empty({n}, mu.options());
    const int64_t need = (int64_t)kp_greedy_workspace_bytes(n, p);
    Tensor ws = workspace.has_value() ? *workspace : at::empty({need < 256 ? 256 : need}, rc.options().dtype(at::kByte));
    if (!(ws.is_cuda() && ws.device() == t.dev && ws.scalar_type() == at::kByte && ws.is_contiguous() && ws.numel() >= need))
        fail("workspace: expected a contiguous uint8 tensor on %s with at least %lld bytes", t.dev.str().c_str(), (long long)need);
    const c10::cuda::CUDAGuard guard(t.dev);
    check(kp_greedy_place(&t.nodes, &t.pods, (int)strategy, (int)filter_mode, out_idx.data_ptr<int64_t>(), out_score.data_ptr<double>(),
                          out_pc.data_ptr<int64_t>(), out_cu.data_ptr<double>(), out_mu.data_ptr<double>(), ws.data_ptr(), (size_t)ws.numel(),
                          c10::cuda::getCurrentCUDAStream(t.dev.index()).stream()),
          "kp_greedy_place");
    return {out_idx, out_score, out_pc, out_cu, out_mu};
}

std::tuple<Tensor, Tensor, Tensor, Tensor> node_scores(const Tensor &ac, const Tensor &am, const Tensor &pc, const Tensor &mp,
                                                       const Tensor &ready, const Tensor &cu, const Tensor &mu, int64_t strategy,
                                                       int64_t filter_mode)
{
    const kp_node_table t = node_table(ac, am, pc, mp, ready, cu, mu);
    const int64_t n = t.n_nodes;
    Tensor score = at::empty({n}, cu.options()), fc = at::empty({n}, cu.options()), fm = at::empty({n}, cu.options());
    Tensor el = at::empty({n}, ready.options());
    const c10::cuda::CUDAGuard guard(pc.device());
    check(kp_node_scores(&t, (int)strategy, (int)filter_mode, score.data_ptr<double>(), fc.data_ptr<double>(), fm.data_ptr<double>(),
                         el.data_ptr<uint8_t>(), c10::cuda::getCurrentCUDAStream(pc.device().index()).stream()),
          "kp_node_scores");
    return {score, fc, fm, el};
}

int64_t snapshot_workspace_bytes(int64_t n_nodes, int64_t n_pods, bool per_pod)
{
    return (int64_t)kp_snapshot_workspace_bytes(n_nodes, n_pods, per_pod ? 1 : 0);
}

int64_t greedy_workspace_bytes(int64_t n_nodes, int64_t n_pods) { return (int64_t)kp_greedy_workspace_bytes(n_nodes, n_pods); }

// the same decisions with the first-feasible probe in front of the table scan (kp_snapshot_place_pruned; an option)
std::tuple<Tensor, Tensor> snapshot_place_pruned(const Tensor &ac, const Tensor &am, const Tensor &pc, const Tensor &mp, const Tensor &ready,
                                                 const Tensor &cu, const Tensor &mu, const Tensor &rc, const Tensor &rm,
                                                 const c10::optional<Tensor> &sid, int64_t strategy, int64_t filter_mode, int64_t probe_depth)
{
    const Tables t = tables(ac, am, pc, mp, ready, cu, mu, rc, rm, sid);
    const int64_t p = rc.numel();
    Tensor out_idx = at::empty({p}, rc.options().dtype(at::kLong));
    Tensor out_score = at::empty({p}, rc.options().dtype(at::kDouble));
    const int64_t ws = (int64_t)kp_snapshot_workspace_bytes(pc.numel(), p, 1);      // the grouping pass needs the pod permutation
    Tensor workspace = at::empty({ws < 256 ? 256 : ws}, rc.options().dtype(at::kByte));
    if (probe_depth < 0 || probe_depth > 1 << 20) fail("probe_depth %lld out of range", (long long)probe_depth);
    const c10::cuda::CUDAGuard guard(t.dev);
    check(kp_snapshot_place_pruned(&t.nodes, &t.pods, (int)strategy, (int)filter_mode, (int)probe_depth, out_idx.data_ptr<int64_t>(),
                                   out_score.data_ptr<double>(), workspace.data_ptr(), (size_t)workspace.numel(),
                                   c10::cuda::getCurrentCUDAStream(t.dev.index()).stream()),
          "kp_snapshot_place_pruned");
    return {out_idx, out_score};
}

}  // namespace

TORCH_LIBRARY(k8s_place, m)
{
    m.def("snapshot_place(Tensor alloc_cpu_milli, Tensor alloc_mem_bytes, Tensor pod_count, Tensor max_pods, Tensor ready, "
          "Tensor cpu_usage_pct, Tensor mem_usage_pct, Tensor req_cpu_milli, Tensor req_mem_bytes, Tensor? strategy_ids, "
          "int strategy, int filter_mode) -> (Tensor, Tensor)");
    m.def("snapshot_place_out(Tensor alloc_cpu_milli, Tensor alloc_mem_bytes, Tensor pod_count, Tensor max_pods, Tensor ready, "
          "Tensor cpu_usage_pct, Tensor mem_usage_pct, Tensor req_cpu_milli, Tensor req_mem_bytes, Tensor? strategy_ids, "
          "int strategy, int filter_mode, Tensor(a!) out_idx, Tensor(b!) out_score, Tensor(c!) workspace) -> ()");
    m.def("snapshot_place_pruned(Tensor alloc_cpu_milli, Tensor alloc_mem_bytes, Tensor pod_count, Tensor max_pods, Tensor ready, "
          "Tensor cpu_usage_pct, Tensor mem_usage_pct, Tensor req_cpu_milli, Tensor req_mem_bytes, Tensor? strategy_ids, "
          "int strategy, int filter_mode, int probe_depth) -> (Tensor, Tensor)");
    m.def("greedy_place(Tensor alloc_cpu_milli, Tensor alloc_mem_bytes, Tensor pod_count, Tensor max_pods, Tensor ready, "
          "Tensor cpu_usage_pct, Tensor mem_usage_pct, Tensor req_cpu_milli, Tensor req_mem_bytes, Tensor? strategy_ids, "
          "int strategy, int filter_mode, Tensor? workspace) -> (Tensor, Tensor, Tensor, Tensor, Tensor)");
    m.def("node_scores(Tensor alloc_cpu_milli, Tensor alloc_mem_bytes, Tensor pod_count, Tensor max_pods, Tensor ready, "
          "Tensor cpu_usage_pct, Tensor mem_usage_pct, int strategy, int filter_mode) -> (Tensor, Tensor, Tensor, Tensor)");
    m.def("snapshot_workspace_bytes(int n_nodes, int n_pods, bool per_pod) -> int", &snapshot_workspace_bytes);
    m.def("greedy_workspace_bytes(int n_nodes, int n_pods) -> int", &greedy_workspace_bytes);
}

TORCH_LIBRARY_IMPL(k8s_place, CUDA, m)
{
    m.impl("snapshot_place", &snapshot_place);
    m.impl("snapshot_place_out", &snapshot_place_out);
    m.impl("snapshot_place_pruned", &snapshot_place_pruned);
    m.impl("greedy_place", &greedy_place);
    m.impl("node_scores", &node_scores);
}
