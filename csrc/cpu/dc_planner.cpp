// Profile-guided scheduler of the ZeRO-3 unit executor ("DeepCompile" counterpart).
//
// The reference rewrites a traced FX graph: zero3_compile inserts allgather/release nodes, the prefetch pass moves the
// all-gathers as early as the profiled memory allows (compile/passes/prefetch.py:41-186), selective_gather keeps the
// parameters with the highest gather time per byte resident inside the profiled head-room
// (compile/passes/selective_gather.py:77-225), and csrc/compile/z3.cpp executes the result. Here there is no traced
// graph: the engine already runs at flat-unit granularity on explicit streams, so the same decisions are a *plan*
// over the unit timeline (forward uses, then backward uses), computed natively from one profiled step:
//
//   * which units stay resident for good (time per byte, inside the head-room below the memory budget),
//   * which units stay gathered between two uses (forward -> backward) instead of being released and re-gathered,
//   * at which step each remaining gather is launched: latest-start scheduling on a serial communication stream
//     against the profiled compute timeline (with a safety factor), moved later where the memory budget or the
//     in-flight cap would be exceeded,
//   * a two-stream simulation of the plan and of the default "one unit ahead" policy (time, exposed wait, peak).
//
// The plan is deterministic in its inputs: every rank feeds the all-reduced profile and gets the same collective order.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <vector>

namespace {

struct Use {
    int step;       // timeline position of the use
    int unit;
    int launch;     // step at whose start the gather is issued (== step: no look-ahead), -1: no gather (kept/persistent)
    int release;    // last step the gathered buffer is alive for this use (step itself, or the next use when kept)
    bool kept_from_prev;
};

struct Sim {
    double total_us, exposed_us;
    int64_t peak_bytes;
};

inline double comm_us(double lat, double bpu, int64_t bytes) { return lat + (bpu > 0 ? (double)bytes / bpu : 0.0); }

// Two-stream simulation: compute steps run back to back, gathers run serially on one communication stream from the
// moment the step that issues them starts, a step waits for the gather of its own unit.
Sim simulate(int n_steps, const double* step_us, const int64_t* step_peak, const std::vector<Use>& uses,
             const int64_t* unit_bytes, int64_t resident_bytes, double lat, double bpu)
{
    std::vector<std::vector<int>> at(n_steps);
    std::vector<int> use_of_step(n_steps, -1);
    for (int i = 0; i < (int)uses.size(); i++) {
        use_of_step[uses[i].step] = i;
        if (uses[i].launch >= 0) at[uses[i].launch].push_back(i);
    }
    for (auto& v : at) std::sort(v.begin(), v.end(), [&](int a, int b) { return uses[a].step < uses[b].step; });
    std::vector<double> ready(uses.size(), 0.0);
    std::vector<int64_t> live(n_steps + 1, 0);
    for (auto& u : uses) {
        if (u.launch < 0) continue;
        for (int t = u.launch; t <= u.release; t++) live[t] += unit_bytes[u.unit];
    }
    double now = 0, comm_free = 0, exposed = 0;
    int64_t peak = 0;
    for (int s = 0; s < n_steps; s++) {
        for (int i : at[s]) {
            const double st = std::max(comm_free, now);
            comm_free = st + comm_us(lat, bpu, unit_bytes[uses[i].unit]);
            ready[i] = comm_free;
        }
        const int i = use_of_step[s];
        if (i >= 0 && uses[i].launch >= 0 && ready[i] > now) {
            exposed += ready[i] - now;
            now = ready[i];
        }
        now += step_us[s];
        peak = std::max(peak, step_peak[s] + live[s] + resident_bytes);
    }
    return {now, exposed, peak};
}

}  // namespace

extern "C" {

// Least-squares fit of time = latency + bytes / bandwidth over measured gathers. out[0] latency (us), out[1] bytes/us.
int dsb200_dc_fit_comm(int n, const int64_t* bytes, const double* us, double* out)
{
    if (n <= 0) return -1;
    double sx = 0, sy = 0, sxx = 0, sxy = 0;
    for (int i = 0; i < n; i++) {
        const double x = (double)bytes[i];
        sx += x, sy += us[i], sxx += x * x, sxy += x * us[i];
    }
    const double den = n * sxx - sx * sx;
    double slope = den > 0 ? (n * sxy - sx * sy) / den : 0.0;
    double lat = (sy - slope * sx) / n;
    if (slope <= 0 || lat < 0) {  // degenerate sample (one size, noise): pure bandwidth model through the mean
        lat = 0;
        slope = sx > 0 ? sy / sx : 0;
    }
    out[0] = lat;
    out[1] = slope > 0 ? 1.0 / slope : 0.0;
    return 0;
}

// Plans one training step.
//   units : unit_bytes[n_units] gathered size; persistent_in[n_units] != 0 -> already resident (never gathered)
//   steps : step_unit[n_steps] unit used by the step (-1 none); step_us[] compute time; step_peak[] peak device bytes
//           inside the step NOT counting gathered non-persistent unit buffers
//   comm  : latency_us + bytes / bytes_per_us, one serial stream
//   limits: mem_budget (bytes; <= 0 unlimited), max_inflight (gathered-ahead bytes not yet used; <= 0 unlimited),
//           persist_headroom (bytes that may be spent on new resident units; < 0: mem_budget - max(step_peak) - largest
//           two units), reuse_window_us (keep a unit between two uses at most this far apart; < 0: 2x its gather time),
//           safety (>= 1: launch early enough for safety x the predicted communication time)
// outputs: persist_out[n_units]; launch_step[n_steps] (-1: no gather at this use), keep_after[n_steps] (1: the unit is
//          not released after this step), sim[8] = {plan total us, plan exposed us, plan peak bytes, default total us,
//          default exposed us, default peak bytes, persisted bytes, feasible (1/0)}
int dsb200_dc_plan(int n_units, const int64_t* unit_bytes, const uint8_t* persistent_in, int n_steps,
                   const int* step_unit, const double* step_us, const int64_t* step_peak, double latency_us,
                   double bytes_per_us, int64_t mem_budget, int64_t max_inflight, int64_t persist_headroom,
                   double reuse_window_us, double safety, uint8_t* persist_out, int* launch_step, int* keep_after,
                   double* sim)
{
    if (n_units <= 0 || n_steps <= 0) return -1;
    for (int s = 0; s < n_steps; s++)
        if (step_unit[s] >= n_units) return -2;
    if (safety < 1) safety = 1;
    const bool limited = mem_budget > 0;
    std::vector<uint8_t> persist(persistent_in, persistent_in + n_units);
    std::vector<int> n_uses(n_units, 0);
    for (int s = 0; s < n_steps; s++)
        if (step_unit[s] >= 0) n_uses[step_unit[s]]++;

    // ---- 1. selective persistence -----------------------------------------------------------------------------
    int64_t max_peak = 0;
    for (int s = 0; s < n_steps; s++) max_peak = std::max(max_peak, step_peak[s]);
    std::vector<int64_t> sorted_b;
    for (int u = 0; u < n_units; u++)
        if (!persist[u] && n_uses[u]) sorted_b.push_back(unit_bytes[u]);
    std::sort(sorted_b.rbegin(), sorted_b.rend());
    int64_t working = 0;  // the unit in use + one in flight must always fit next to the resident set
    for (size_t i = 0; i < sorted_b.size() && i < 2; i++) working += sorted_b[i];
    int64_t headroom = persist_headroom;
    if (headroom < 0) headroom = limited ? std::max<int64_t>(0, mem_budget - max_peak - working) : 0;
    std::vector<int> cand;
    for (int u = 0; u < n_units; u++)
        if (!persist[u] && n_uses[u] && unit_bytes[u] > 0) cand.push_back(u);
    // gather time per byte, all uses counted: latency-dominated (small) units first
    std::stable_sort(cand.begin(), cand.end(), [&](int a, int b) {
        const double ta = n_uses[a] * comm_us(latency_us, bytes_per_us, unit_bytes[a]) / (double)unit_bytes[a];
        const double tb = n_uses[b] * comm_us(latency_us, bytes_per_us, unit_bytes[b]) / (double)unit_bytes[b];
        return ta > tb;
    });
    int64_t persisted = 0;
    for (int u : cand) {
        if (persisted + unit_bytes[u] > headroom) break;
        persisted += unit_bytes[u];
        persist[u] = 1;
    }
    std::memcpy(persist_out, persist.data(), n_units);

    // ---- 2. uses, keep-between-uses ---------------------------------------------------------------------------
    std::vector<double> t0(n_steps + 1, 0.0);  // ideal compute timeline
    for (int s = 0; s < n_steps; s++) t0[s + 1] = t0[s] + step_us[s];
    std::vector<Use> uses;
    std::vector<int> last_use(n_units, -1);
    std::vector<int64_t> extra(n_steps, 0);  // gathered bytes alive at each step under the plan so far
    auto fits = [&](int from, int to, int64_t bytes) {
        if (!limited) return true;
        for (int t = from; t <= to; t++)
            if (step_peak[t] + persisted + extra[t] + bytes > mem_budget) return false;
        return true;
    };
    for (int s = 0; s < n_steps; s++) {
        launch_step[s] = -1;
        keep_after[s] = 0;
        const int u = step_unit[s];
        if (u < 0 || persist[u]) continue;
        Use x{s, u, s, s, false};
        const int p = last_use[u];
        if (p >= 0) {
            Use& prev = uses[p];
            const double gap = t0[s] - t0[prev.step + 1];
            const double win = reuse_window_us >= 0 ? reuse_window_us
                                                    : 2.0 * comm_us(latency_us, bytes_per_us, unit_bytes[u]);
            if (gap <= win && fits(prev.step + 1, s, unit_bytes[u])) {
                for (int t = prev.step + 1; t <= s; t++) extra[t] += unit_bytes[u];
                prev.release = s;
                keep_after[prev.step] = 1;
                x.launch = -1;
                x.kept_from_prev = true;
            }
        }
        if (x.launch >= 0) extra[s] += unit_bytes[u];
        last_use[u] = (int)uses.size();
        uses.push_back(x);
    }

    // ---- 3. launch steps: latest-start scheduling in reverse, then the memory / in-flight limits ---------------
    bool feasible = true;
    double next_start = 1e300;  // start time of the following job on the communication stream
    std::vector<int64_t> inflight(n_steps, 0);  // gathered-ahead bytes (launched, not yet at their use step)
    for (int i = (int)uses.size() - 1; i >= 0; i--) {
        Use& x = uses[i];
        if (x.launch < 0) continue;
        const int64_t b = unit_bytes[x.unit];
        const double dur = comm_us(latency_us, bytes_per_us, b) * safety;
        const double finish = std::min(t0[x.step], next_start);
        const double start = finish - dur;
        next_start = start;
        int e = x.step;  // the step whose start is the launch point: the last one starting at or before `start`
        while (e > 0 && t0[e] > start) e--;
        // never launch before the previous use of the same unit is over (it would still be gathered or in release)
        for (int j = i - 1; j >= 0; j--)
            if (uses[j].unit == x.unit) {
                e = std::max(e, uses[j].release + 1);
                break;
            }
        e = std::min(e, x.step);
        // memory: the buffer exists from the launch step on
        int lo = x.step;
        while (lo > e) {
            const int t = lo - 1;
            if (limited && step_peak[t] + persisted + extra[t] + b > mem_budget) break;
            if (max_inflight > 0 && inflight[t] + b > max_inflight) break;
            lo--;
        }
        x.launch = lo;
        for (int t = lo; t < x.step; t++) {
            extra[t] += b;
            inflight[t] += b;
        }
        if (limited && step_peak[x.step] + persisted + extra[x.step] > mem_budget) feasible = false;
    }
    // The stream is FIFO: a gather issued before one that is needed earlier would delay it. Make the launch points
    // non-decreasing in use order (moving a launch later only shortens the buffer's life, so the limits still hold).
    int prev_launch = 0;
    for (auto& x : uses) {
        if (x.launch < 0) continue;
        x.launch = std::min(x.step, std::max(x.launch, prev_launch));
        prev_launch = x.launch;
        launch_step[x.step] = x.launch;
    }

    // ---- 4. simulation of the plan and of the default policy ---------------------------------------------------
    const Sim sp = simulate(n_steps, step_us, step_peak, uses, unit_bytes, persisted, latency_us, bytes_per_us);
    std::vector<Use> dflt;
    {
        std::vector<int> pos_of;  // steps that use a non-resident unit (input residency only)
        for (int s = 0; s < n_steps; s++)
            if (step_unit[s] >= 0 && !persistent_in[step_unit[s]]) pos_of.push_back(s);
        for (size_t k = 0; k < pos_of.size(); k++) {
            const int s = pos_of[k];
            int launch = k == 0 ? s : pos_of[k - 1];  // gathered when the previous unit starts
            if (k > 0 && step_unit[pos_of[k - 1]] == step_unit[s]) launch = s;  // same unit back to back: re-gather
            dflt.push_back({s, step_unit[s], launch, s, false});
        }
    }
    const Sim sd = simulate(n_steps, step_us, step_peak, dflt, unit_bytes, 0, latency_us, bytes_per_us);
    sim[0] = sp.total_us, sim[1] = sp.exposed_us, sim[2] = (double)sp.peak_bytes;
    sim[3] = sd.total_us, sim[4] = sd.exposed_us, sim[5] = (double)sd.peak_bytes;
    sim[6] = (double)persisted, sim[7] = feasible ? 1.0 : 0.0;
    return 0;
}

}  // extern "C"
