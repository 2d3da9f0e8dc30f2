/*
 * CPU oracle (C restatement) for the fallback placement path.
 * TEST INFRASTRUCTURE, NOT PRODUCT CODE: only tests/, __graft_entry__.smoke() and
 * bench.py's CPU-baseline / reference legs may load the library built from this file.
 *
 * It restates, over the int64/float64 wire tables, the same arithmetic as
 * oracle/py_oracle.py (which restates /root/reference/scheduler.py and is pinned to
 * golden vectors produced by the reference itself).  It exists because the CPython
 * oracle needs ~0.6 us per pod x node pair: this one lets parity tests run at
 * 10^8 pairs in seconds.  tests/test_oracle_golden.py checks it bit-for-bit against the
 * CPython oracle and against the JSON files under tests/golden.  Parity status: PINNED (via those).
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -shared -fPIC (see oracle/Makefile).
 * -ffp-contract=off matters: a fused multiply-add changes score bits.
 *
 * Brute force on purpose: the score is recomputed for every pod x node pair, exactly
 * like the reference's loop (scheduler.py:531-549).  No hoisting, no sorting.
 */
#include <stdint.h>
#include <stddef.h>

#define KPO_FILTER_READY 0
#define KPO_FILTER_CAPACITY 1

typedef struct {
    int64_t n;
    const int64_t *alloc_cpu_milli;
    const int64_t *alloc_mem_bytes;
    const int64_t *pod_count;
    const int64_t *max_pods;
    const uint8_t *ready;
    const double *cpu_usage;
    const double *mem_usage;
} kpo_nodes;

/* scheduler.py:150-151 */
static double usage_a3(int64_t pc, int64_t mp)
{
    return mp > 0 ? ((double)pc / (double)mp) * 50.0 : 0.0;
}

/* scheduler.py:537-541, left-associated sum */
static double score_rb(double cu, double mu, int64_t pc, int64_t mp)
{
    double cpu_score = (100.0 - cu) / 100.0;
    double mem_score = (100.0 - mu) / 100.0;
    double pod_score = mp > 0 ? (double)(mp - pc) / (double)mp : 0.0;
    double a = cpu_score * 0.35;
    double b = mem_score * 0.35;
    double c = pod_score * 0.30;
    double ab = a + b;
    return ab + c;
}

/* scheduler.py:542-543 */
static double score_ll(double cu, double mu)
{
    double a = 100.0 - cu;
    double b = 100.0 - mu;
    return a + b;
}

/* scheduler.py:232-233 with available_* rebuilt from the integer wire format
 * (cores = milli/1000.0; GiB = bytes/1024/1024/1024) */
static int pod_fits(const kpo_nodes *t, int64_t i, double cu, double mu, int64_t pc,
                    double cpu_req, double mem_req)
{
    if (!(pc < t->max_pods[i])) return 0;
    double avail_cpu = (double)t->alloc_cpu_milli[i] / 1000.0;
    double avail_mem = (double)t->alloc_mem_bytes[i] / 1024.0 / 1024.0 / 1024.0;
    double free_cpu = avail_cpu * (100.0 - cu) / 100.0;
    double free_mem = avail_mem * (100.0 - mu) / 100.0;
    return cpu_req <= free_cpu && mem_req <= free_mem;
}

/* filter -> score -> first-best argmax for ONE pod (scheduler.py:528-549).
 * pc/cu/mu may point at mutable copies (greedy). */
static int64_t pick(const kpo_nodes *t, const int64_t *pc, const double *cu, const double *mu,
                    int strategy, int filter_mode, double cpu_req, double mem_req,
                    double *out_score)
{
    int64_t best = -1;
    if (strategy == 0 || strategy == 1) {
        double best_score = -1.0;
        for (int64_t i = 0; i < t->n; ++i) {
            if (!t->ready[i]) continue;
            if (filter_mode == KPO_FILTER_CAPACITY &&
                !pod_fits(t, i, cu[i], mu[i], pc[i], cpu_req, mem_req)) continue;
            double s = strategy == 0 ? score_rb(cu[i], mu[i], pc[i], t->max_pods[i])
                                     : score_ll(cu[i], mu[i]);
            if (s > best_score) { best_score = s; best = i; }
        }
        *out_score = best_score;
    } else {
        int64_t best_score = -1;            /* integer compare, like Python ints */
        for (int64_t i = 0; i < t->n; ++i) {
            if (!t->ready[i]) continue;
            if (filter_mode == KPO_FILTER_CAPACITY &&
                !pod_fits(t, i, cu[i], mu[i], pc[i], cpu_req, mem_req)) continue;
            int64_t s = pc[i];
            if (s > best_score) { best_score = s; best = i; }
        }
        *out_score = (double)best_score;
    }
    return best;
}

static double req_cores(int64_t milli) { return (double)milli / 1000.0; }
static double req_gib(int64_t bytes) { return (double)bytes / 1024.0 / 1024.0 / 1024.0; }

/* Snapshot: pods [p0, p1) against an unchanging table.  strategy_ids may be NULL
 * (then `strategy` applies to all).  Ids other than 0/1 take the round_robin branch. */
int kpo_snapshot_range(int64_t n_nodes, const int64_t *alloc_cpu_milli,
                       const int64_t *alloc_mem_bytes, const int64_t *pod_count,
                       const int64_t *max_pods, const uint8_t *ready,
                       const double *cpu_usage, const double *mem_usage,
                       int64_t p0, int64_t p1, const int64_t *req_cpu_milli,
                       const int64_t *req_mem_bytes, const int64_t *strategy_ids,
                       int strategy, int filter_mode, int64_t *out_idx, double *out_score)
{
    kpo_nodes t = { n_nodes, alloc_cpu_milli, alloc_mem_bytes, pod_count, max_pods, ready,
                    cpu_usage, mem_usage };
    for (int64_t j = p0; j < p1; ++j) {
        int s = strategy_ids ? (strategy_ids[j] == 0 ? 0 : strategy_ids[j] == 1 ? 1 : 2)
                             : (strategy == 0 ? 0 : strategy == 1 ? 1 : 2);
        double sc;
        out_idx[j] = pick(&t, pod_count, cpu_usage, mem_usage, s, filter_mode,
                          req_cores(req_cpu_milli[j]), req_gib(req_mem_bytes[j]), &sc);
        out_score[j] = sc;
    }
    return 0;
}

/* Greedy: pods in order; each placement commits (scheduler.py:598 -> :144-151):
 * pod_count += 1 and both usages re-derived.  pod_count / cpu_usage / mem_usage are
 * updated IN PLACE. */
int kpo_greedy(int64_t n_nodes, const int64_t *alloc_cpu_milli, const int64_t *alloc_mem_bytes,
               int64_t *pod_count, const int64_t *max_pods, const uint8_t *ready,
               double *cpu_usage, double *mem_usage,
               int64_t n_pods, const int64_t *req_cpu_milli, const int64_t *req_mem_bytes,
               const int64_t *strategy_ids, int strategy, int filter_mode,
               int64_t *out_idx, double *out_score)
{
    kpo_nodes t = { n_nodes, alloc_cpu_milli, alloc_mem_bytes, pod_count, max_pods, ready,
                    cpu_usage, mem_usage };
    for (int64_t j = 0; j < n_pods; ++j) {
        int s = strategy_ids ? (strategy_ids[j] == 0 ? 0 : strategy_ids[j] == 1 ? 1 : 2)
                             : (strategy == 0 ? 0 : strategy == 1 ? 1 : 2);
        double sc;
        int64_t i = pick(&t, pod_count, cpu_usage, mem_usage, s, filter_mode,
                         req_cores(req_cpu_milli[j]), req_gib(req_mem_bytes[j]), &sc);
        out_idx[j] = i;
        out_score[j] = sc;
        if (i >= 0) {
            pod_count[i] += 1;
            double u = usage_a3(pod_count[i], max_pods[i]);
            cpu_usage[i] = u;
            mem_usage[i] = u;
        }
    }
    return 0;
}

/* single-node helpers for known-answer tests */
double kpo_score(int strategy, double cu, double mu, int64_t pc, int64_t mp)
{
    return strategy == 0 ? score_rb(cu, mu, pc, mp) : strategy == 1 ? score_ll(cu, mu) : (double)pc;
}
double kpo_usage(int64_t pc, int64_t mp) { return usage_a3(pc, mp); }
int kpo_abi_version(void) { return 1; }
