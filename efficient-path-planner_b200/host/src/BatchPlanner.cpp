// BatchPlanner.cpp — see BatchPlanner.h.
#include "epp/BatchPlanner.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <limits>
#include <queue>
#include <unordered_map>

namespace epp {

namespace {

uint64_t mix64(uint64_t seed, uint64_t counter)
{
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (counter + 1);
    z ^= z >> 30;
    z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27;
    z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

double u01(uint64_t seed, uint64_t index, unsigned lane)
{
    return static_cast<double>(mix64(seed, index * 8 + lane) >> 11) * 0x1.0p-53;
}

double dist(const double *a, const double *b)
{
    const double dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    return std::sqrt(dx * dx + dy * dy + dz * dz);
}

bool bit(const std::vector<uint32_t> &m, std::size_t i) { return (m[i >> 5] >> (i & 31)) & 1u; }

double now()
{
    using namespace std::chrono;
    return duration<double>(steady_clock::now().time_since_epoch()).count();
}

} // namespace

bool BatchPlanner::solve(const Vector3d &start, const Vector3d &goal, double timeLimit, std::vector<Vector3d> &path,
                         BatchPlannerStats *stats) const
{
    BatchPlannerStats local;
    BatchPlannerStats &st = stats ? *stats : local;
    st = BatchPlannerStats();
    const double t0 = now();
    path.clear();

    // start / goal validity in one batch of two (si->isValid on start and goal in the OMPL planners)
    {
        const double sg[6] = {start(0), start(1), start(2), goal(0), goal(1), goal(2)};
        uint32_t w = 0;
        sv_.isValidBatch(sg, 2, &w);
        if ((w & 3u) != 3u) {
            st.seconds = now() - t0;
            return false;
        }
    }

    const double ext[3] = {hi_(0) - lo_(0), hi_(1) - lo_(1), hi_(2) - lo_(2)};
    const double volume = ext[0] * ext[1] * ext[2];
    std::size_t n = initialSamples_;
    for (;; n *= 2) {
        ++st.rounds;
        // 1. samples: index 0 = start, 1 = goal, n uniform in the bounds, then n/8 in boxes of half-width range and
        //    range/4 around start and goal (a checkpoint sits 0.21 m behind its gate, often in a slot between the
        //    gate and an obstacle that uniform samples reach only after several doublings)
        const std::size_t nt = std::max<std::size_t>(64, n / 8) & ~std::size_t(3);
        const std::size_t total = n + 2 + nt;
        std::vector<double> xyz(3 * total);
        for (int d = 0; d < 3; ++d) {
            xyz[d] = start(d);
            xyz[3 + d] = goal(d);
        }
        for (std::size_t i = 0; i < n; ++i)
            for (unsigned d = 0; d < 3; ++d) xyz[3 * (i + 2) + d] = lo_(d) + u01(seed_ + st.rounds, i, d) * ext[d];
        for (std::size_t j = 0; j < nt; ++j) {
            const Vector3d &c = (j & 1) ? goal : start;
            const double half = (j & 2) ? 0.25 * range_ : range_;
            for (unsigned d = 0; d < 3; ++d) {
                const double x = c(d) + (2.0 * u01(seed_ + st.rounds, n + j, d) - 1.0) * half;
                xyz[3 * (n + 2 + j) + d] = std::min(std::max(x, lo_(d)), hi_(d));
            }
        }
        std::vector<uint32_t> vmask(World::maskWords(total));
        sv_.isValidBatch(xyz.data(), total, vmask.data());
        std::vector<double> pts;
        pts.reserve(3 * total);
        for (std::size_t i = 0; i < total; ++i)
            if (bit(vmask, i)) pts.insert(pts.end(), xyz.begin() + 3 * i, xyz.begin() + 3 * i + 3);
        const std::size_t m = pts.size() / 3;
        st.samples += n;
        st.validSamples += m;

        // 2. radius neighbours through a hash grid of cell size r
        const double gamma = 2.0 * std::cbrt((1.0 + 1.0 / 3.0) * volume / (4.0 / 3.0 * M_PI));
        const double r = std::max(range_ * std::cbrt(2048.0 / static_cast<double>(n)),
                                  gamma * std::cbrt(std::log(static_cast<double>(m)) / static_cast<double>(m)));
        auto cellOf = [&](const double *p, int d) { return static_cast<int64_t>(std::floor((p[d] - lo_(d)) / r)); };
        auto key = [](int64_t x, int64_t y, int64_t z) { return (x * 73856093) ^ (y * 19349663) ^ (z * 83492791); };
        std::unordered_map<int64_t, std::vector<uint32_t>> grid;
        grid.reserve(m);
        for (std::size_t i = 0; i < m; ++i)
            grid[key(cellOf(&pts[3 * i], 0), cellOf(&pts[3 * i], 1), cellOf(&pts[3 * i], 2))].push_back((uint32_t)i);
        std::vector<uint32_t> ea, eb;
        for (std::size_t i = 0; i < m; ++i) {
            const double *p = &pts[3 * i];
            const int64_t cx = cellOf(p, 0), cy = cellOf(p, 1), cz = cellOf(p, 2);
            for (int64_t dx = -1; dx <= 1; ++dx)
                for (int64_t dy = -1; dy <= 1; ++dy)
                    for (int64_t dz = -1; dz <= 1; ++dz) {
                        auto it = grid.find(key(cx + dx, cy + dy, cz + dz));
                        if (it == grid.end()) continue;
                        for (uint32_t j : it->second)
                            if (j > i && dist(p, &pts[3 * j]) <= r) {
                                ea.push_back((uint32_t)i);
                                eb.push_back(j);
                            }
                    }
        }
        const std::size_t ne = ea.size();
        std::vector<double> A(3 * ne), B(3 * ne);
        for (std::size_t e = 0; e < ne; ++e)
            for (int d = 0; d < 3; ++d) {
                A[3 * e + d] = pts[3 * ea[e] + d];
                B[3 * e + d] = pts[3 * eb[e] + d];
            }
        std::vector<uint32_t> emask(World::maskWords(ne));
        if (ne) mv_.checkMotionBatch(A.data(), B.data(), ne, emask.data());
        st.edgesChecked += ne;

        // 3. Dijkstra start (0) -> goal (1) over the valid edges
        std::vector<std::vector<std::pair<uint32_t, double>>> adj(m);
        for (std::size_t e = 0; e < ne; ++e)
            if (bit(emask, e)) {
                const double w = dist(&pts[3 * ea[e]], &pts[3 * eb[e]]);
                adj[ea[e]].push_back({eb[e], w});
                adj[eb[e]].push_back({ea[e], w});
                ++st.validEdges;
            }
        std::vector<double> cost(m, std::numeric_limits<double>::infinity());
        std::vector<int64_t> parent(m, -1);
        using QE = std::pair<double, uint32_t>;
        std::priority_queue<QE, std::vector<QE>, std::greater<QE>> pq;
        cost[0] = 0.0;
        pq.push({0.0, 0});
        while (!pq.empty()) {
            const QE top = pq.top();
            pq.pop();
            if (top.first > cost[top.second]) continue;
            if (top.second == 1) break;
            for (const auto &nb : adj[top.second]) {
                const double c = top.first + nb.second;
                if (c < cost[nb.first]) {
                    cost[nb.first] = c;
                    parent[nb.first] = top.second;
                    pq.push({c, nb.first});
                }
            }
        }
        if (std::isfinite(cost[1])) {
            std::vector<Vector3d> raw;
            for (int64_t v = 1; v >= 0; v = parent[v]) raw.push_back(Vector3d(pts[3 * v], pts[3 * v + 1], pts[3 * v + 2]));
            std::reverse(raw.begin(), raw.end());
            path = reduceVertices(raw, &st);
            st.pathLength = 0.0;
            for (std::size_t i = 1; i < path.size(); ++i) st.pathLength += dist(path[i - 1].data(), path[i].data());
            st.seconds = now() - t0;
            return true;
        }
        if (now() - t0 > timeLimit || 2 * n > maxSamples_) break;
    }
    st.seconds = now() - t0;
    return false;
}

std::vector<Vector3d> BatchPlanner::reduceVertices(const std::vector<Vector3d> &path, BatchPlannerStats *stats) const
{
    const std::size_t k = path.size();
    if (k < 3) return path;
    // every shortcut (i, j), j > i + 1, in one edge batch
    std::vector<double> A, B;
    std::vector<std::pair<uint32_t, uint32_t>> ij;
    for (uint32_t i = 0; i + 2 < k; ++i)
        for (uint32_t j = i + 2; j < k; ++j) {
            ij.push_back({i, j});
            for (int d = 0; d < 3; ++d) {
                A.push_back(path[i](d));
                B.push_back(path[j](d));
            }
        }
    std::vector<uint32_t> mask(World::maskWords(ij.size()));
    mv_.checkMotionBatch(A.data(), B.data(), ij.size(), mask.data());
    if (stats) stats->shortcutChecks += ij.size();
    // shortest chain 0 -> k-1 over consecutive edges (known valid) and valid shortcuts
    std::vector<double> cost(k, std::numeric_limits<double>::infinity());
    std::vector<int64_t> parent(k, -1);
    cost[0] = 0.0;
    std::vector<std::vector<uint32_t>> into(k);
    for (std::size_t e = 0; e < ij.size(); ++e)
        if (bit(mask, e)) into[ij[e].second].push_back(ij[e].first);
    for (uint32_t j = 1; j < k; ++j) {
        into[j].push_back(j - 1);
        for (uint32_t i : into[j]) {
            const double c = cost[i] + dist(path[i].data(), path[j].data());
            if (c < cost[j]) {
                cost[j] = c;
                parent[j] = i;
            }
        }
    }
    std::vector<Vector3d> out;
    for (int64_t v = (int64_t)k - 1; v >= 0; v = parent[v]) out.push_back(path[v]);
    std::reverse(out.begin(), out.end());
    return out;
}

} // namespace epp
