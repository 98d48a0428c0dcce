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

// Profile mode.  The reference's rule gives device i a share proportional to range_i / time_i: it treats the cost of a
// work-item as a property of the DEVICE.  When the cost depends on the POSITION in the range (Mandelbrot rows), a device
// that is handed more items gets its neighbour's items at the neighbour's cost, the linear extrapolation overshoots, and
// the split oscillates (8 B200s, round 1: 16 calls to settle; one slow device among 8 never settles in simulation).
// Here the measured times are read as a piecewise-constant cost density over the global range,
//     density(x) = (time_i + 0.01) / range_i   for x in device i's segment,
// and the new boundaries cut the integral of that density into equal parts; every call refines the density around the
// current boundaries.  The move toward those boundaries is damped (0.8; 0.5 if the imbalance grew since the last call,
// which is what a device-speed difference looks like to a position model; 0.3 once max/mean time <= 1.05 so timing noise
// does not jitter a settled split).  Same snap-to-step and sum fix-up as the reference.
static void profileBalance(const std::vector<double>& benchmark, BalanceState& st, int totalRange, std::vector<int>& ranges,
                           int step) {
    const int n = (int)ranges.size();
    std::vector<double> len, dens;   // active segments in range order
    double total = 0.0, mx = 0.0;
    for (int i = 0; i < n; i++)
        if (ranges[i] > 0) {
            const double t = benchmark[i] + 0.01;
            len.push_back((double)ranges[i]);
            dens.push_back(t / (double)ranges[i]);
            total += t;
            mx = t > mx ? t : mx;
        }
    if (len.empty()) return;
    const double imbalance = mx * (double)len.size() / total;
    const double target = total / (double)n;
    std::vector<double> want(n, 0.0);
    size_t seg = 0;
    double used = 0.0, assigned = 0.0;
    for (int d = 0; d + 1 < n; d++) {
        double need = target, length = 0.0;
        while (seg < len.size()) {
            const double avail = (len[seg] - used) * dens[seg];
            if (avail >= need - 1e-15) {
                const double take = dens[seg] > 0.0 ? need / dens[seg] : 0.0;
                used += take;
                length += take;
                break;
            }
            need -= avail;
            length += len[seg] - used;
            seg++;
            used = 0.0;
        }
        want[d] = length;
        assigned += length;
    }
    want[n - 1] = (double)totalRange - assigned;

    double gain = imbalance > 1.05 ? 0.8 : 0.3;
    if (st.lastImbalance > 0.0 && imbalance > 1.02 * st.lastImbalance && imbalance > 1.05) gain = 0.5;
    st.lastImbalance = imbalance;

    for (int i = 0; i < n; i++) {
        const int tmp = (int)((double)ranges[i] + gain * (want[i] - (double)ranges[i]));
        const int rem = tmp % step;
        ranges[i] = (rem < step / 2) ? tmp - rem : tmp + (step - rem);
        if (ranges[i] < 0) ranges[i] = 0;
    }
    while (sum(ranges) > totalRange) ranges[maxIndex(ranges)] -= step;
    while (sum(ranges) < totalRange) ranges[maxIndex(ranges)] += step;
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

    if (mode == BalancerMode::Profile) {
        profileBalance(benchmark, st, totalRange, ranges, step);
        return;
    }

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
