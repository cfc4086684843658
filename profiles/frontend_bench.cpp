// Router fan-out through the batching front-end, as BASELINE.json configs[4] states it: 512 tickers x 256 episodes,
// the child actors (here: host threads) queue their SelectionAction and UpdateQ messages, one qp_flush per fan-out
// answers all selects with one launch and turns all updates into one gradient-SUM step.
// Build + run: profiles/frontend_bench.sh   (g++ against include/qpolicy.h, links libqpolicy.so at run time)
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

#include "qpolicy.h"

static void check(int st, const char* what) {
    if (st != 0) {
        std::fprintf(stderr, "%s: status %d: %s\n", what, st, qp_last_error());
        std::exit(1);
    }
}

int main(int argc, char** argv) {
    const int64_t n_msgs = argc > 1 ? std::atoll(argv[1]) : 512 * 256;
    const int n_threads = argc > 2 ? std::atoi(argv[2]) : 16;
    const int fanouts = argc > 3 ? std::atoi(argv[3]) : 6;
    qp_config cfg;
    check(qp_config_default(&cfg), "qp_config_default");
    qp_handle* h = nullptr;
    check(qp_create(&cfg, &h), "qp_create");
    const int64_t D = qp_input_dim(h);
    // one state per episode: a quarter-tick window around 100, budget, shares (values do not matter for the timing)
    std::vector<float> states((size_t)n_msgs * D), next((size_t)n_msgs * D);
    for (int64_t e = 0; e < n_msgs; ++e) {
        for (int64_t k = 0; k < D - 2; ++k) {
            states[(size_t)e * D + k] = 100.0f + 0.25f * (float)((e * 7 + k * 3) % 41);
            next[(size_t)e * D + k] = 100.0f + 0.25f * (float)((e * 7 + (k + 1) * 3) % 41);
        }
        states[(size_t)e * D + D - 2] = next[(size_t)e * D + D - 2] = 1000.0f;
        states[(size_t)e * D + D - 1] = next[(size_t)e * D + D - 1] = 1.0f;
    }
    std::vector<int64_t> tickets((size_t)n_msgs);
    std::vector<int32_t> actions((size_t)n_msgs);
    double t_enq = 0, t_flush = 0, t_take = 0, t_total = 0;
    for (int f = 0; f < fanouts; ++f) {
        const auto t0 = std::chrono::steady_clock::now();
        {   // SelectionAction from every child, on n_threads dispatcher threads
            std::vector<std::thread> th;
            for (int t = 0; t < n_threads; ++t)
                th.emplace_back([&, t] {
                    for (int64_t e = t; e < n_msgs; e += n_threads)
                        check(qp_enqueue_select(h, &states[(size_t)e * D], D, (float)(500 + f), e, &tickets[(size_t)e]), "enqueue_select");
                });
            for (auto& x : th) x.join();
        }
        const auto t1 = std::chrono::steady_clock::now();
        int64_t ns = 0, nu = 0;
        check(qp_flush(h, &ns, &nu), "qp_flush (selects)");
        const auto t2 = std::chrono::steady_clock::now();
        for (int64_t e = 0; e < n_msgs; ++e) check(qp_take_action(h, tickets[(size_t)e], &actions[(size_t)e]), "take_action");
        const auto t3 = std::chrono::steady_clock::now();
        {   // UpdateQ from every child
            std::vector<std::thread> th;
            for (int t = 0; t < n_threads; ++t)
                th.emplace_back([&, t] {
                    for (int64_t e = t; e < n_msgs; e += n_threads)
                        check(qp_enqueue_update(h, &states[(size_t)e * D], D, 0.25f * (float)(actions[(size_t)e] - 1),
                                                &next[(size_t)e * D], D), "enqueue_update");
                });
            for (auto& x : th) x.join();
        }
        const auto t4 = std::chrono::steady_clock::now();
        check(qp_flush(h, &ns, &nu), "qp_flush (updates)");
        const auto t5 = std::chrono::steady_clock::now();
        if (f >= 1) {   // the first fan-out warms the queues and the device buffers up
            auto s = [](auto a, auto b) { return std::chrono::duration<double>(b - a).count(); };
            t_enq += s(t0, t1) + s(t3, t4);
            t_flush += s(t1, t2) + s(t4, t5);
            t_take += s(t2, t3);
            t_total += s(t0, t5);
        }
    }
    const int n = fanouts - 1;
    std::printf("front-end, %lld episodes per fan-out, %d host threads: %.2f ms per fan-out (enqueue %.2f, two flushes %.2f, "
                "take_action %.2f) -> %.4g env-steps/s\n",
                (long long)n_msgs, n_threads, t_total / n * 1e3, t_enq / n * 1e3, t_flush / n * 1e3, t_take / n * 1e3,
                (double)n_msgs * n / t_total);
    qp_destroy(h);
    return 0;
}
