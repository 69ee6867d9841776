// Library-wide state: last error text, launch counter, version.
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
