// balance.h — the iterative per-compute-id load balancer (pure host arithmetic, no CUDA).
//
// Follows the reference's Functions.loadBalance and its history helpers
// (HelperFunctions.cs:119-156 and :190-280) with identical double/int arithmetic — the `+0.01`
// guards, the `(int)` truncations, the 0.3 damping, the snap-to-step rule and the two fix-up loops —
// and the first-call rule of Cores.compute (Cores.cs:569-596: equal split, remainder to device 0,
// bench = 10 ms for everyone, then one balancing pass).
#pragma once
#include <vector>

namespace cekirdekler {

constexpr int kHistoryDepth = 10;  // Cores.cs:1065

enum class BalancerMode {
    Reference = 0,  // smoothing gate exactly as HelperFunctions.cs:239 ("window full")
    Adaptive = 1,   // same arithmetic, but the average only spans samples taken while the devices were balanced
                    // (max/mean time <= 1.05): a transient discards the window, so stale shares are never mixed in
                    // (SURVEY.md quirk Q14).  Follows the un-smoothed trajectory while converging, smooths after.
    Profile = 2,    // reads the device times as a cost density over the range and cuts its integral into equal parts
                    // (see profileBalance in balance.cpp): for work whose cost depends on the position in the range
};

struct BalanceState {
    // normalised throughput shares, oldest first: history[k][device]
    std::vector<std::vector<double>> history;
    std::vector<double> lastThroughputs;  // Cores.tmpThroughputs — exposed as normalizedComputePowersOfDevices
    double lastImbalance = 0.0;  // Adaptive mode: max/mean device time seen by the previous call
    int valid = 0;  // Adaptive mode: newest history rows that belong to the current balanced regime
    void reset(int devices);
};

// One balancing pass.  `ranges` is updated in place; Σ ranges == totalRange and every range is a multiple of step.
void loadBalance(const std::vector<double>& benchmark, bool smooth, BalanceState& st, int totalRange,
                 std::vector<int>& ranges, int step, BalancerMode mode = BalancerMode::Reference);

// First call for a compute id: equal split with the remainder on device 0 (Cores.cs:573-580).
void initialSplit(int totalRange, int devices, std::vector<int>& ranges);

// step rule of Cores.cs:595/604
int balanceStep(const std::vector<int>& ranges, bool pipelineEnabled, int blobs, int localRange, int totalRange);

}  // namespace cekirdekler
