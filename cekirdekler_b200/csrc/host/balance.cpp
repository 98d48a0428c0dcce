// balance.cpp — see balance.h.  Reference: HelperFunctions.cs:84-97 (maxIndex), :119-156 (history),
// :190-280 (loadBalance); Cores.cs:569-605 (first call, step rule).
#include "balance.h"

namespace cekirdekler {

void BalanceState::reset(int devices) {
    history.assign(kHistoryDepth, std::vector<double>(devices, 0.0));
    lastThroughputs.assign(devices, 0.0);
    valid = 0;
    lastImbalance = 0.0;
}

static int maxIndex(const std::vector<int>& a) {  // first index holding the maximum (HelperFunctions.cs:84-97)
    int r = 0, v = -100000000;
    for (int i = 0; i < (int)a.size(); i++)
        if (v < a[i]) {
            v = a[i];
            r = i;
        }
    return r;
}

static long long sum(const std::vector<int>& a) {
    long long s = 0;
    for (int v : a) s += v;
    return s;
}

void loadBalance(const std::vector<double>& benchmark, bool smooth, BalanceState& st, int totalRange,
                 std::vector<int>& ranges, int step, BalancerMode mode) {
    const int n = (int)ranges.size();
    if ((int)st.history.size() != kHistoryDepth || (int)st.history[0].size() != n) st.reset(n);
    std::vector<int> tmpRanges(n, 0);
    std::vector<double>& thr = st.lastThroughputs;
    thr.assign(n, 0.0);

    double totalBenchmark = 0.0;
    for (int i = 0; i < n; i++) totalBenchmark += benchmark[i];
    totalBenchmark += 0.01 * n;
    double totalThroughput = 0.0;
    for (int i = 0; i < n; i++) {
        thr[i] = (totalBenchmark / (benchmark[i] + 0.01)) * (ranges[i] + 1);
        totalThroughput += thr[i];
    }
    if (totalThroughput <= 0.0000001) totalThroughput = 0.01;

    std::vector<double> normalized(n, 0.0);
    if (smooth) {
        for (int i = 0; i < n; i++) normalized[i] = thr[i] / totalThroughput;
        // shift one step older, newest last (performanceHistoryShiftOld)
        for (int k = 0; k + 1 < kHistoryDepth; k++) st.history[k] = st.history[k + 1];
        st.history[kHistoryDepth - 1] = normalized;
        std::vector<double> avg(n, 0.0);
        if (mode == BalancerMode::Adaptive) {
            // window = newest rows recorded while max/mean device time stayed within 5 %
            double mx = 0.0, sm = 0.0;
            int act = 0;
            for (int i = 0; i < n; i++)
                if (ranges[i] > 0) {
                    mx = benchmark[i] > mx ? benchmark[i] : mx;
                    sm += benchmark[i];
                    act++;
                }
            const bool balanced = act > 0 && sm > 0.0 && mx * act <= 1.05 * sm;
            st.valid = balanced ? (st.valid < kHistoryDepth ? st.valid + 1 : kHistoryDepth) : 1;
            for (int k = kHistoryDepth - st.valid; k < kHistoryDepth; k++)
                for (int j = 0; j < n; j++) avg[j] += st.history[k][j];
            for (int j = 0; j < n; j++) avg[j] /= (double)st.valid;
        } else {
            // column mean over the whole window (performanceHistoryAverage)
            for (int k = 0; k < kHistoryDepth; k++)
                for (int j = 0; j < n; j++) avg[j] += st.history[k][j];
            const double div = 1.0f / ((double)kHistoryDepth);
            for (int j = 0; j < n; j++) avg[j] *= div;
        }
        normalized = avg;
    }

    // Step size toward the target split.  The reference always moves 30 % of the way (HelperFunctions.cs:247).  Adaptive
    // mode moves further while the devices are far from balance — gain = clamp(2 (max/mean - 1), 0.3, 0.6) — and uses the
    // reference's 0.3 near balance and whenever the imbalance grew since the previous call.  (Larger gains oscillate on
    // non-uniform workloads: with timing noise a 0.9 cap diverged on 8 GPUs in round 1, in the simulation and on the box.)
    double gain = 0.3;
    if (mode == BalancerMode::Adaptive) {
        double mx = 0.0, sm = 0.0;
        int act = 0;
        for (int i = 0; i < n; i++)
            if (ranges[i] > 0) {
                mx = benchmark[i] > mx ? benchmark[i] : mx;
                sm += benchmark[i];
                act++;
            }
        if (act > 0 && sm > 0.0) {
            const double imbalance = mx * act / sm;
            const double g = 2.0 * (imbalance - 1.0);
            gain = g < 0.3 ? 0.3 : (g > 0.6 ? 0.6 : g);
            if (st.lastImbalance > 0.0 && imbalance > 1.02 * st.lastImbalance) gain = 0.3;
            st.lastImbalance = imbalance;
        }
    }

    for (int i = 0; i < n; i++) {
        const bool useAverage = smooth && (mode == BalancerMode::Adaptive || st.history[0][0] > 0.00001);
        const double share = useAverage ? normalized[i] : (thr[i] / totalThroughput);
        if (ranges[i] != 0) {
            tmpRanges[i] = ranges[i] - (int)((ranges[i] - (totalRange * share)) * gain);
        } else {
            tmpRanges[i] = (int)(((totalRange * (totalBenchmark / (benchmark[i] + 0.01)) / totalThroughput)));
        }
    }

    for (int i = 0; i < n; i++) {
        const int rem = tmpRanges[i] % step;
        if (rem < step / 2) ranges[i] = tmpRanges[i] - rem;
        else ranges[i] = tmpRanges[i] + (step - rem);
    }
    while (sum(ranges) > totalRange) ranges[maxIndex(ranges)] -= step;
    while (sum(ranges) < totalRange) ranges[maxIndex(ranges)] += step;
}

void initialSplit(int totalRange, int devices, std::vector<int>& ranges) {
    ranges.assign(devices, totalRange / devices);
    long long s = 0;
    for (int v : ranges) s += v;
    if (s != totalRange) ranges[0] += (int)(totalRange - s);
}

int balanceStep(const std::vector<int>& ranges, bool pipelineEnabled, int blobs, int localRange, int totalRange) {
    bool b1 = true;
    for (int v : ranges) b1 = b1 && (v >= blobs * localRange);
    return (b1 && pipelineEnabled && (totalRange >= blobs * localRange)) ? blobs * localRange : localRange;
}

}  // namespace cekirdekler
