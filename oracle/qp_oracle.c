/*
 * qp_oracle.c — CPU restatement of the reference's Q-learning decision policy hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it, and
 * only as the checker or as the timed CPU baseline.  The CUDA library never links or calls it.
 *
 * PARITY UNPINNED.  The reference (cosmir17/Scala-akka-tensorflow-sharetrade-helper) cannot be
 * executed here (no JVM/sbt; its TensorFlow artefact org.platanios:tensorflow:0.4.1 is pinned to
 * darwin-cpu-x86_64, build.sbt:21) and its own tests assert message types only — no Q-value,
 * loss, weight or reward (src/test/scala/QDecisionPolicyActorSpec.scala:27-45).  This file
 * therefore restates the algorithm from the reference's call sites, plus the published semantics
 * of the TensorFlow ops that the graph at QDecisionPolicyActor.scala:38-50 instantiates:
 *   MatMul/Add/Relu/Sub/Square, gradients of those, and ResourceApplyAdagrad
 *   (accum += g*g; var -= g*lr*rsqrt(accum)), AdaGrad(0.01f) initial accumulator taken on trust
 *   as 1e-8 (configurable), RandomNormalInitializer() = N(0,1), argmax = first maximal index.
 * It is cross-checked in tests/ against an independent autograd statement of the same graph.
 *
 * Compile twice: -DQPO_REAL=float (the oracle proper; -ffp-contract=off, fixed summation
 * order) and -DQPO_REAL=double (the fp64 twin, used only to calibrate the fp32 tolerance).
 *
 * Reference lines followed (all relative to the reference root):
 *   hyper-parameters ........ src/main/scala/QDecisionPolicyActor.scala:17-22, :50
 *   network forward ......... QDecisionPolicyActor.scala:38-47
 *   loss / optimiser ........ QDecisionPolicyActor.scala:49-50
 *   SelectionAction ......... QDecisionPolicyActor.scala:56-62
 *   UpdateQ ................. QDecisionPolicyActor.scala:64-76, :85-89
 *   episode loop ............ src/main/scala/TrainerChildActor.scala:64-71, :82-103
 *   decision book-keeping ... TrainerChildActor.scala:118-123
 *   reward / next state ..... TrainerChildActor.scala:136-146, :154-160
 *   portfolio statistics .... src/main/scala/TrainerRouterActor.scala:149-151
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

#ifndef QPO_REAL
#define QPO_REAL float
#endif
typedef QPO_REAL real;

#define QPO_ACTIONS 3 /* Buy=0, Sell=1, Hold=2 — QDecisionPolicyActor.scala:17 */

typedef struct {
    int32_t wn;          /* price-window length: reference 201 (TrainerChildActor.scala:90) */
    int32_t hidden;      /* h1Dim = 200 (QDecisionPolicyActor.scala:22) */
    float epsilon;       /* 0.9f   (:19) */
    float gamma;         /* 0.001f (:20) */
    float lr;            /* AdaGrad(0.01f) (:50) */
    float bias;          /* constant 0.1f biases (:42, :46) */
    float step_div;      /* step / 1000f (:58) */
    int32_t bookkeeping; /* 0 = REFERENCE (decision base = constructor position, TrainerChildActor.scala:118-123)
                            1 = RLPY      (decision base = folded carry, as in rl.py) */
    uint64_t seed;       /* Philox key */
} qpo_config;

/* ------------------------------------------------------------------------------------------
 * Philox4x32-10 (Salmon et al., Random123).  The reference uses the unseeded global
 * scala.util.Random (QDecisionPolicyActor.scala:58, :62); a counter-based generator replaces it
 * so that CPU and GPU draw identical numbers.  Draw discipline kept from the reference: one
 * 24-bit uniform per select (nextFloat), one bounded integer only when exploring (nextInt(3)).
 * ------------------------------------------------------------------------------------------ */
void qpo_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
    uint32_t k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static void philox_at(uint64_t seed, uint64_t id, uint32_t t, uint32_t stream, uint32_t out[4]) {
    uint32_t ctr[4] = {(uint32_t)id, (uint32_t)(id >> 32), t, stream};
    uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    qpo_philox4x32_10(ctr, key, out);
}

/* Synthetic quarter-tick random walk (stream 0): exactly representable in fp32, no libm. */
void qpo_gen_prices(uint64_t seed, int64_t first_series, int64_t n_series, int32_t T, float* out) {
    for (int64_t s = 0; s < n_series; ++s) {
        float* p = out + s * (int64_t)T;
        float v = 100.0f;
        p[0] = v;
        for (int32_t t = 0; t + 1 < T; ++t) {
            uint32_t r[4];
            philox_at(seed, (uint64_t)(first_series + s), (uint32_t)t, 0u, r);
            int k = (int)(r[0] % 5u);
            v = v + 0.25f * (float)(k - 2);
            if (v < 1.0f) v = 1.0f;
            p[t + 1] = v;
        }
    }
}

/* Host-side N(0,1) initialiser (stream 2), Box-Muller on two 24-bit uniforms.  Stands in for the
 * unseeded tf.RandomNormalInitializer() at QDecisionPolicyActor.scala:41, :45. */
void qpo_gen_normal(uint64_t seed, int64_t first, int64_t n, float* out) {
    for (int64_t i = 0; i < n; ++i) {
        uint32_t r[4];
        philox_at(seed, (uint64_t)(first + i), 0u, 2u, r);
        double u1 = ((double)(r[0] >> 8) + 0.5) * (1.0 / 16777216.0);
        double u2 = ((double)(r[1] >> 8) + 0.5) * (1.0 / 16777216.0);
        out[i] = (float)(sqrt(-2.0 * log(u1)) * cos(6.283185307179586476925286766559 * u2));
    }
}

/* ------------------------------------------------------------------------------------------
 * Network forward — QDecisionPolicyActor.scala:43, :47.  h = relu(x·W1 + b1), q = relu(h·W2 + b2);
 * note the ReLU on the output layer.  W1 is [D,H] row-major, W2 is [H,3] row-major.
 * Each output element is accumulated sequentially over k = 0..D-1 in `real` with no FMA
 * contraction; the k-outer loop order below gives exactly that per-element order.
 * ------------------------------------------------------------------------------------------ */
static void forward(const qpo_config* c, const real* W1, const real* W2, const real* x,
                    real* z, real* h, real* q) {
    const int D = c->wn + 2, H = c->hidden;
    for (int j = 0; j < H; ++j) z[j] = (real)0;
    for (int k = 0; k < D; ++k) {
        const real xk = x[k];
        const real* w = W1 + (size_t)k * H;
        for (int j = 0; j < H; ++j) z[j] = z[j] + xk * w[j];
    }
    const real b = (real)c->bias;
    for (int j = 0; j < H; ++j) {
        real v = z[j] + b;
        h[j] = v > (real)0 ? v : (real)0;
    }
    for (int a = 0; a < QPO_ACTIONS; ++a) {
        real acc = (real)0;
        for (int j = 0; j < H; ++j) acc = acc + h[j] * W2[(size_t)j * QPO_ACTIONS + a];
        acc = acc + b;
        q[a] = acc > (real)0 ? acc : (real)0;
    }
}

void qpo_forward(const qpo_config* c, const real* params, const real* x, real* h_out, real* q_out) {
    const int D = c->wn + 2, H = c->hidden;
    real* z = (real*)malloc(sizeof(real) * (size_t)H);
    forward(c, params, params + (size_t)D * H, x, z, h_out, q_out);
    free(z);
}

/* argmax with first-max-wins ties, as Tensor.argmax (QDecisionPolicyActor.scala:60, :69). */
static int argmax3(const real* q) {
    int k = 0;
    if (q[1] > q[k]) k = 1;
    if (q[2] > q[k]) k = 2;
    return k;
}

/* SelectionAction — QDecisionPolicyActor.scala:56-62.  Returns the action index; *explored is
 * set when the random branch was taken.  q must already hold q(state). */
static int select_action(const qpo_config* c, const real* q, float step, uint64_t env_id,
                         uint32_t step_ctr, int* explored) {
    uint32_t r[4];
    philox_at(c->seed, env_id, step_ctr, 1u, r);
    const float u = (float)(r[0] >> 8) * (1.0f / 16777216.0f);
    float thr = step / c->step_div;
    if (c->epsilon < thr) thr = c->epsilon;
    if (u < thr) {
        *explored = 0;
        return argmax3(q);
    }
    *explored = 1;
    return (int)(((uint64_t)r[1] * 3u) >> 32);
}

int qpo_select_action(const qpo_config* c, const real* params, const real* x, float step,
                      uint64_t env_id, real* q_out, int* explored) {
    const int D = c->wn + 2, H = c->hidden;
    real* buf = (real*)malloc(sizeof(real) * (size_t)H * 2);
    forward(c, params, params + (size_t)D * H, x, buf, buf + H, q_out);
    free(buf);
    int ex = 0;
    int a = select_action(c, q_out, step, env_id, (uint32_t)step, &ex);
    if (explored) *explored = ex;
    return a;
}

/* UpdateQ gradient — QDecisionPolicyActor.scala:64-73 with the graph's loss at :49-50.
 *   k = argmax q(s'); target = reward + gamma*q(s')[k]; y = q(s) with y[k] = target
 *   loss = sum((y - q(s))^2)  ->  only column k has a non-zero gradient.
 * Accumulates d loss / d params into g (same layout as params).  z/h scratch: 4*H reals. */
typedef struct {
    real q_s[QPO_ACTIONS];
    real q_n[QPO_ACTIONS];
    int32_t k_next;
    real target;
} qpo_td_info;

static void td_grad(const qpo_config* c, const real* params, const real* s, real reward,
                    const real* s2, real* g, real* scratch, qpo_td_info* info, int force_k) {
    const int D = c->wn + 2, H = c->hidden;
    const real* W1 = params;
    const real* W2 = params + (size_t)D * H;
    real* z = scratch;
    real* h = scratch + H;
    real* h2 = scratch + 2 * H;
    real* dz1 = scratch + 3 * H;
    real q[QPO_ACTIONS], qn[QPO_ACTIONS];
    forward(c, W1, W2, s, z, h, q);
    forward(c, W1, W2, s2, z, h2, qn);
    /* force_k >= 0: a test imposes the index (used where the top-two gap of q(s') is inside the
     * stated tolerance and the implementation under test legitimately chose the other one) */
    const int k = force_k >= 0 ? force_k : argmax3(qn);
    const real target = reward + (real)c->gamma * qn[k];
    /* d/dq (y-q)^2 = -2 (y-q); ReluGrad passes where the output is strictly positive */
    const real dq = -((real)2 * (target - q[k]));
    const real dz2 = q[k] > (real)0 ? dq : (real)0;
    real* gW1 = g;
    real* gW2 = g + (size_t)D * H;
    for (int j = 0; j < H; ++j) {
        gW2[(size_t)j * QPO_ACTIONS + k] = gW2[(size_t)j * QPO_ACTIONS + k] + h[j] * dz2;
        const real dh = dz2 * W2[(size_t)j * QPO_ACTIONS + k];
        dz1[j] = h[j] > (real)0 ? dh : (real)0;
    }
    for (int r = 0; r < D; ++r) {
        const real xr = s[r];
        real* row = gW1 + (size_t)r * H;
        for (int j = 0; j < H; ++j) row[j] = row[j] + xr * dz1[j];
    }
    if (info) {
        for (int a = 0; a < QPO_ACTIONS; ++a) { info->q_s[a] = q[a]; info->q_n[a] = qn[a]; }
        info->k_next = k;
        info->target = target;
    }
}

/* ResourceApplyAdagrad: accum += g*g; var -= g*lr*rsqrt(accum)  (no epsilon in the denominator). */
void qpo_adagrad(const qpo_config* c, real* params, real* acc, const real* g, int64_t n) {
    const real lr = (real)c->lr;
    for (int64_t i = 0; i < n; ++i) {
        const real gi = g[i];
        const real a = acc[i] + gi * gi;
        acc[i] = a;
        params[i] = params[i] - (gi * lr) * ((real)1 / (real)sqrt((double)a));
    }
}

/* Gradient of a batch of explicit (s, reward, s') rows, summed in row order (batched-SUM rule:
 * the graph's x placeholder is Shape(-1, inputDim), QDecisionPolicyActor.scala:38, and minimize
 * of the un-reduced loss sums over every element). */
void qpo_td_grad_batch(const qpo_config* c, const real* params, const real* s, const float* reward,
                       const real* s2, int64_t B, real* g, real* q_s, real* q_n, int32_t* k_next,
                       real* target) {
    const int D = c->wn + 2, H = c->hidden;
    const int64_t P = (int64_t)D * H + (int64_t)H * QPO_ACTIONS;
    real* scratch = (real*)malloc(sizeof(real) * (size_t)H * 4);
    memset(g, 0, sizeof(real) * (size_t)P);
    for (int64_t b = 0; b < B; ++b) {
        qpo_td_info info;
        td_grad(c, params, s + b * D, (real)reward[b], s2 + b * D, g, scratch, &info, -1);
        for (int a = 0; a < QPO_ACTIONS; ++a) {
            if (q_s) q_s[b * QPO_ACTIONS + a] = info.q_s[a];
            if (q_n) q_n[b * QPO_ACTIONS + a] = info.q_n[a];
        }
        if (k_next) k_next[b] = info.k_next;
        if (target) target[b] = info.target;
    }
    free(scratch);
}

/* UpdateQ for a batch: gradient SUM then one AdaGrad step.  B = 1 is the reference's message. */
void qpo_update_q(const qpo_config* c, real* params, real* acc, const real* s, const float* reward,
                  const real* s2, int64_t B, real* g_out) {
    const int D = c->wn + 2, H = c->hidden;
    const int64_t P = (int64_t)D * H + (int64_t)H * QPO_ACTIONS;
    real* g = g_out ? g_out : (real*)malloc(sizeof(real) * (size_t)P);
    qpo_td_grad_batch(c, params, s, reward, s2, B, g, NULL, NULL, NULL, NULL);
    qpo_adagrad(c, params, acc, g, P);
    if (!g_out) free(g);
}

/* ------------------------------------------------------------------------------------------
 * Episode step — TrainerChildActor.scala:82-103 for env-step index i of every env.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
    /* all optional (NULL = not recorded); one entry per env */
    int32_t* action_sel;  /* index returned by the policy */
    int32_t* action_eff;  /* after Buy/Sell demotion to Hold (TrainerChildActor.scala:118-123) */
    int32_t* explored;
    int32_t* k_next;
    float* reward;        /* fp64 reward cast to fp32 at TrainerChildActor.scala:158 */
    real* q_s;            /* [n,3] */
    real* q_n;            /* [n,3] */
    real* target;
} qpo_trace;

/* Teacher hook of the parity tests (all optional): entries >= 0 replace the oracle's own choice for
 * that env.  Only used after the unforced run has shown that the differing choice lies inside the
 * stated tolerance, so that gradient, weights and carry can still be compared in every case. */
typedef struct {
    const int32_t* action;   /* replaces the index SelectionAction returns */
    const int32_t* k_next;   /* replaces argmax q(s') in UpdateQ */
} qpo_force;

typedef struct {
    double* budget;       /* carry._1 */
    int32_t* shares;      /* carry._2 */
    double* share_value;  /* carry._3, seeded 0.0 (TrainerChildActor.scala:84-86) */
} qpo_env_state;

static void env_one(const qpo_config* c, const real* params, const float* p, int32_t i,
                    double B0, int32_t N0, int64_t e, uint64_t env_id, qpo_env_state* st,
                    const qpo_trace* tr, real* g, real* scratch, real* s, real* s2, const qpo_force* fc) {
    const int wn = c->wn, D = wn + 2, H = c->hidden;
    const double budget = st->budget[e];
    const int32_t shares = st->shares[e];
    /* state — TrainerChildActor.scala:89-91 */
    for (int k = 0; k < wn; ++k) s[k] = (real)p[i + k];
    s[wn] = (real)(float)budget;
    s[wn + 1] = (real)(float)shares;
    /* current portfolio — :92 */
    const double cur = budget + (double)shares * st->share_value[e];
    /* ask SelectionAction — :93 */
    real q[QPO_ACTIONS];
    forward(c, params, params + (size_t)D * H, s, scratch, scratch + H, q);
    int explored = 0;
    int a = select_action(c, q, (float)i, env_id, (uint32_t)i, &explored);
    if (fc && fc->action && fc->action[e] >= 0) a = fc->action[e];
    /* new share value — :94 */
    const double p_new = (double)p[i + wn];
    /* decision — :118-123 (Q1: REFERENCE mode decides from the constructor's position) */
    const double base_b = c->bookkeeping == 0 ? B0 : budget;
    const int32_t base_n = c->bookkeeping == 0 ? N0 : shares;
    double nb = base_b;
    int32_t ns = base_n;
    int a_eff = 2;
    if (a == 0 && base_b >= p_new) { nb = base_b - p_new; ns = base_n + 1; a_eff = 0; }
    else if (a == 1 && base_n > 0) { nb = base_b + p_new; ns = base_n - 1; a_eff = 1; }
    /* reward and next state — :136-146 */
    const double new_portfolio = nb + (double)ns * p_new;
    const double reward = new_portfolio - cur;
    const float r32 = (float)reward;
    for (int k = 0; k < wn; ++k) s2[k] = (real)p[i + 1 + k];
    s2[wn] = (real)(float)nb;
    s2[wn + 1] = (real)(float)ns;
    /* ask UpdateQ — :97 */
    qpo_td_info info;
    td_grad(c, params, s, (real)r32, s2, g, scratch, &info, (fc && fc->k_next) ? fc->k_next[e] : -1);
    /* carry — :99-102 */
    st->budget[e] = nb;
    st->shares[e] = ns;
    st->share_value[e] = p_new;
    if (tr) {
        if (tr->action_sel) tr->action_sel[e] = a;
        if (tr->action_eff) tr->action_eff[e] = a_eff;
        if (tr->explored) tr->explored[e] = explored;
        if (tr->k_next) tr->k_next[e] = info.k_next;
        if (tr->reward) tr->reward[e] = r32;
        if (tr->target) tr->target[e] = info.target;
        for (int x = 0; x < QPO_ACTIONS; ++x) {
            if (tr->q_s) tr->q_s[e * QPO_ACTIONS + x] = info.q_s[x];
            if (tr->q_n) tr->q_n[e * QPO_ACTIONS + x] = info.q_n[x];
        }
    }
}

/* One synchronous env-step for all envs: every env selects and forms its TD gradient with the
 * same weights, gradients are summed in env order, one AdaGrad step follows.
 * mode 0 = batched SUM (the build's rule), mode 1 = sequential per-sample (the reference's
 * one-mailbox order: env 0 select+update, then env 1 with the updated weights, ...).
 * env_id_base offsets the RNG counter so that a rank's shard draws the global env's numbers. */
void qpo_env_step_forced(const qpo_config* c, real* params, real* acc, const float* prices, int32_t T,
                         const int32_t* series_of_env, int64_t n_envs, int64_t env_id_base, int32_t i,
                         double B0, int32_t N0, qpo_env_state* st, const qpo_trace* tr, int mode,
                         real* g_out, const qpo_force* fc) {
    const int D = c->wn + 2, H = c->hidden;
    const int64_t P = (int64_t)D * H + (int64_t)H * QPO_ACTIONS;
    real* g = g_out ? g_out : (real*)malloc(sizeof(real) * (size_t)P);
    real* scratch = (real*)malloc(sizeof(real) * ((size_t)H * 4 + (size_t)D * 2));
    real* s = scratch + (size_t)H * 4;
    real* s2 = s + D;
    memset(g, 0, sizeof(real) * (size_t)P);
    for (int64_t e = 0; e < n_envs; ++e) {
        const float* p = prices + (int64_t)series_of_env[e] * T;
        env_one(c, params, p, i, B0, N0, e, (uint64_t)(env_id_base + e), st, tr, g, scratch, s, s2, fc);
        if (mode == 1) {
            qpo_adagrad(c, params, acc, g, P);
            memset(g, 0, sizeof(real) * (size_t)P);
        }
    }
    if (mode == 0) qpo_adagrad(c, params, acc, g, P);
    free(scratch);
    if (!g_out) free(g);
}

void qpo_env_step(const qpo_config* c, real* params, real* acc, const float* prices, int32_t T,
                  const int32_t* series_of_env, int64_t n_envs, int64_t env_id_base, int32_t i,
                  double B0, int32_t N0, qpo_env_state* st, const qpo_trace* tr, int mode,
                  real* g_out) {
    qpo_env_step_forced(c, params, acc, prices, T, series_of_env, n_envs, env_id_base, i, B0, N0, st, tr, mode, g_out, NULL);
}

/* Multi-threaded batched-SUM step used only as the timed CPU baseline ("all host cores").
 * Threads own contiguous env ranges; per-thread partial gradients are added in thread order,
 * so the summation order differs from qpo_env_step (results agree within fp32 tolerance). */
typedef struct {
    const qpo_config* c; const real* params; const float* prices; int32_t T;
    const int32_t* series_of_env; int64_t n_envs; int64_t env_id_base; int32_t i;
    double B0; int32_t N0; qpo_env_state* st; real* partials; int t; int nt;
} mt_job;

static void* mt_worker(void* arg) {
    mt_job* j = (mt_job*)arg;
    const qpo_config* c = j->c;
    const int D = c->wn + 2, H = c->hidden;
    const int64_t P = (int64_t)D * H + (int64_t)H * QPO_ACTIONS;
    real* g = j->partials + (size_t)j->t * P;
    real* scratch = (real*)malloc(sizeof(real) * ((size_t)H * 4 + (size_t)D * 2));
    real* s = scratch + (size_t)H * 4;
    real* s2 = s + D;
    memset(g, 0, sizeof(real) * (size_t)P);
    const int64_t lo = j->n_envs * j->t / j->nt, hi = j->n_envs * (j->t + 1) / j->nt;
    for (int64_t e = lo; e < hi; ++e) {
        const float* p = j->prices + (int64_t)j->series_of_env[e] * j->T;
        env_one(c, j->params, p, j->i, j->B0, j->N0, e, (uint64_t)(j->env_id_base + e), j->st, NULL,
                g, scratch, s, s2, NULL);
    }
    free(scratch);
    return NULL;
}

void qpo_env_step_mt(const qpo_config* c, real* params, real* acc, const float* prices, int32_t T,
                     const int32_t* series_of_env, int64_t n_envs, int64_t env_id_base, int32_t i,
                     double B0, int32_t N0, qpo_env_state* st, int n_threads, real* partials) {
    const int D = c->wn + 2, H = c->hidden;
    const int64_t P = (int64_t)D * H + (int64_t)H * QPO_ACTIONS;
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    pthread_t th[256];
    mt_job jobs[256];
    for (int t = 0; t < n_threads; ++t) {
        mt_job j = {c, params, prices, T, series_of_env, n_envs, env_id_base, i, B0, N0, st, partials, t, n_threads};
        jobs[t] = j;
        if (t > 0) pthread_create(&th[t], NULL, mt_worker, &jobs[t]);
    }
    mt_worker(&jobs[0]);
    for (int t = 1; t < n_threads; ++t) pthread_join(th[t], NULL);
    for (int u = 1; u < n_threads; ++u) {
        const real* src = partials + (size_t)u * P;
        for (int64_t k = 0; k < P; ++k) partials[k] = partials[k] + src[k];
    }
    qpo_adagrad(c, params, acc, partials, P);
}

int qpo_max_threads(void) {
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n < 1 ? 1 : (int)n;
}

/* Router statistics — TrainerRouterActor.scala:149-151: mean and POPULATION standard deviation. */
void qpo_portfolio_stats(const double* x, int64_t n, double* mean_out, double* std_out) {
    double sum = 0.0;
    for (int64_t i = 0; i < n; ++i) sum += x[i];
    const double mean = sum / (double)n;
    double v = 0.0;
    for (int64_t i = 0; i < n; ++i) v += pow(x[i] - mean, 2.0);
    *mean_out = mean;
    *std_out = sqrt(v / (double)n);
}

int qpo_sizeof_real(void) { return (int)sizeof(real); }
