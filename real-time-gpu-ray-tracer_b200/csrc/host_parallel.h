// host_parallel.h — fn(k) for k in [0, n) on up to 8 host threads (contiguous chunks); serial when n is small.
// RTB200_HOST_THREADS caps the thread count (several ranks / renderers sharing one host: cores / (ranks x renderers)).
#pragma once

#include <algorithm>
#include <cstddef>
#include <cstdlib>
#include <thread>
#include <vector>

template<class F>
static void parallel_for(size_t n, bool parallel, F fn, size_t min_per_thread = 1) {
    static const unsigned hw = std::thread::hardware_concurrency();
    static const size_t cap = getenv("RTB200_HOST_THREADS") ? (size_t) std::max(1, atoi(getenv("RTB200_HOST_THREADS"))) : 8u;
    size_t T = parallel ? std::min<size_t>(std::min<size_t>(hw ? hw : 1u, cap), n / std::max<size_t>(1, min_per_thread)) : 1;
    if (T <= 1) { for (size_t k = 0; k < n; k++) fn(k); return; }
    std::vector<std::thread> th;
    for (size_t t = 0; t < T; t++)
        th.emplace_back([=]() { for (size_t k = n * t / T; k < n * (t + 1) / T; k++) fn(k); });
    for (std::thread &x: th) x.join();
}
