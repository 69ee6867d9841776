// Library-wide state: last error text, launch counter, version.
#include <stdlib.h>

#include <atomic>
#include <mutex>
#include <string>

#include "../include/arbius_b200.h"

static std::mutex g_err_mu;
static std::string g_err = "";
static std::atomic<unsigned long long> g_launches{0};

extern "C" void ab_set_error(const char* msg) {
    std::lock_guard<std::mutex> lk(g_err_mu);
    g_err = msg ? msg : "";
}
extern "C" const char* ab_last_error(void) {
    std::lock_guard<std::mutex> lk(g_err_mu);
    static thread_local std::string copy;
    copy = g_err;
    return copy.c_str();
}
extern "C" void ab_count_launch(int n) { g_launches.fetch_add(static_cast<unsigned long long>(n), std::memory_order_relaxed); }
extern "C" unsigned long long ab_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
extern "C" int ab_version(void) { return 100; }

// Programmatic dependent launch switch: off unless AB_PDL=1 in the environment; ab_set_pdl() overrides (A/B timing).
// Measured on the 20-step AnythingV3 task inside one CUDA graph: 137.2 ms without, 139.5 ms with — graph edges
// already hide the launch latency and early-resident dependents only compete for shared memory / TMEM.
static std::atomic<int> g_pdl{-1};
extern "C" int ab_pdl_enabled(void) {
    int v = g_pdl.load(std::memory_order_relaxed);
    if (v < 0) {
        const char* e = getenv("AB_PDL");
        v = (e && e[0] == '1') ? 1 : 0;
        g_pdl.store(v, std::memory_order_relaxed);
    }
    return v;
}
extern "C" void ab_set_pdl(int on) { g_pdl.store(on ? 1 : 0, std::memory_order_relaxed); }
