// dev tool: greedy forward search of a brick library for one footprint.
// Bricks are built by doubling: B_k = B_a U (B_b + (du, dv)). The score of a library is the total size of
// the greedy exact covers of the footprint's offset sets (one per heading bin, reference float expression).
//   g++ -O3 -fopenmp -std=c++17 tools/lib_search.cpp -o /tmp/lib_search && /tmp/lib_search rect 20
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <set>
#include <string>
#include <vector>

using Offsets = std::vector<std::pair<int, int>>;
struct Mask
{
    uint64_t w[3] = {0, 0, 0};
    int count() const { return __builtin_popcountll(w[0]) + __builtin_popcountll(w[1]) + __builtin_popcountll(w[2]); }
    bool empty() const { return (w[0] | w[1] | w[2]) == 0; }
    int overlap(const Mask &o) const
    {
        return __builtin_popcountll(w[0] & o.w[0]) + __builtin_popcountll(w[1] & o.w[1]) + __builtin_popcountll(w[2] & o.w[2]);
    }
    void andnot(const Mask &o)
    {
        w[0] &= ~o.w[0];
        w[1] &= ~o.w[1];
        w[2] &= ~o.w[2];
    }
    bool operator<(const Mask &o) const { return memcmp(w, o.w, sizeof(w)) < 0; }
};

static std::vector<std::pair<int, int>> footprint(const std::string &kind)
{
    std::vector<std::pair<int, int>> p;
    auto rim = [&](int x0, int y0, int x1, int y1) {
        for (int x = x0; x <= x1; ++x)
            for (int y = y0; y <= y1; ++y)
                if (x == x0 || x == x1 || y == y0 || y == y1)
                    p.emplace_back(x, y);
    };
    if (kind == "rect")
        rim(-7, -5, 7, 5);
    else if (kind == "box")
        rim(-4, -5, 12, 5);
    else if (kind == "solid")
        for (int x = -7; x <= 7; ++x)
            for (int y = -5; y <= 5; ++y)
                p.emplace_back(x, y);
    return p;
}

static Offsets offsets_for(const std::vector<std::pair<int, int>> &fp, int t, int nt)
{
    const float bin = 2.0f * (float)M_PI / (float)nt;
    const float th = (float)t * bin;
    const float c = cosf(th), s = sinf(th);
    Offsets S;
    const float x = 1000.5f, y = 1000.5f;
    for (auto &q : fp)
    {
        const float px = (float)q.first, py = (float)q.second;
        float a = px * c, b = py * s, u = x + a, xn = u - b;
        float a2 = px * s, b2 = py * c, u2 = y + a2, yn = u2 + b2;
        S.emplace_back((int)floorf(xn) - 1000, (int)floorf(yn) - 1000);
    }
    std::sort(S.begin(), S.end());
    S.erase(std::unique(S.begin(), S.end()), S.end());
    return S;
}

struct SetInfo
{
    Offsets S;
    std::map<std::pair<int, int>, int> rank;
};

static Offsets normalise(Offsets p)
{
    std::sort(p.begin(), p.end());
    p.erase(std::unique(p.begin(), p.end()), p.end());
    const int ox = p[0].first, oy = p[0].second;
    for (auto &q : p)
    {
        q.first -= ox;
        q.second -= oy;
    }
    return p;
}

// placements of brick (normalised, first point at (0,0)) inside set
static void placements(const SetInfo &si, const Offsets &brick, std::vector<Mask> &out)
{
    for (auto &anchor : si.S)
    {
        Mask m;
        bool ok = true;
        for (auto &q : brick)
        {
            auto it = si.rank.find({anchor.first + q.first, anchor.second + q.second});
            if (it == si.rank.end())
            {
                ok = false;
                break;
            }
            m.w[it->second >> 6] |= 1ull << (it->second & 63);
        }
        if (ok)
            out.push_back(m);
    }
}

static int greedy(const std::vector<Mask> &cands, int n)
{
    Mask unc;
    for (int b = 0; b < n; ++b)
        unc.w[b >> 6] |= 1ull << (b & 63);
    int used = 0;
    while (!unc.empty())
    {
        int best = -1, gain = 0;
        for (size_t c = 0; c < cands.size(); ++c)
        {
            const int g = cands[c].overlap(unc);
            if (g > gain)
            {
                gain = g;
                best = (int)c;
            }
        }
        unc.andnot(cands[best]);
        ++used;
    }
    return used;
}

int main(int argc, char **argv)
{
    const std::string kind = argc > 1 ? argv[1] : "rect";
    const int K = argc > 2 ? atoi(argv[2]) : 16;
    const int nt = 72;
    auto fp = footprint(kind);
    std::vector<SetInfo> sets;
    {
        std::set<Offsets> seen;
        for (int t = 0; t < nt; ++t)
        {
            Offsets S = offsets_for(fp, t, nt);
            if (!seen.insert(S).second)
                continue;
            SetInfo si;
            si.S = S;
            for (size_t i = 0; i < S.size(); ++i)
                si.rank[S[i]] = (int)i;
            sets.push_back(si);
        }
    }
    printf("%s: %zu distinct offset sets\n", kind.c_str(), sets.size());
    struct Brick
    {
        Offsets pts;
        int a, b, du, dv;
    };
    std::vector<Brick> lib{{{{0, 0}}, 0, 0, 0, 0}};
    if (argc > 3 && std::string(argv[3]) == "stairs")
    {
        // seed: the STAIRS library of se2_fast.cu (12 tables)
        const int R[11][4] = {{0, 0, 1, 0}, {0, 0, 0, 1}, {0, 0, 1, 1}, {0, 0, 1, -1}, {1, 1, 2, 1}, {1, 1, 2, -1}, {2, 2, 1, 2}, {2, 2, -1, 2}, {1, 1, 2, 0}, {2, 2, 0, 2}, {9, 9, 4, 0}};
        for (auto &r : R)
        {
            Offsets p = lib[r[0]].pts;
            for (auto &q : lib[r[1]].pts)
                p.emplace_back(q.first + r[2], q.second + r[3]);
            std::sort(p.begin(), p.end());
            p.erase(std::unique(p.begin(), p.end()), p.end());
            lib.push_back({p, r[0], r[1], r[2], r[3]});  // NOT normalised: recipes refer to table a's origin
        }
    }
    std::vector<std::vector<Mask>> cur(sets.size());  // placements of the current library per set
    for (size_t s = 0; s < sets.size(); ++s)
        for (auto &b : lib)
            placements(sets[s], normalise(b.pts), cur[s]);
    auto total = [&](const std::vector<std::vector<Mask>> &c) {
        long tot = 0;
        for (size_t s = 0; s < sets.size(); ++s)
            tot += greedy(c[s], (int)sets[s].S.size());
        return tot;
    };
    printf("K=%zu total %ld (%.2f per set)\n", lib.size(), total(cur), (double)total(cur) / sets.size());
    while ((int)lib.size() < K)
    {
        // candidates
        std::map<Offsets, Brick> cands;
        for (size_t a = 0; a < lib.size(); ++a)
            for (size_t b = a; b < lib.size(); ++b)
                for (int du = -12; du <= 12; ++du)
                    for (int dv = -12; dv <= 12; ++dv)
                    {
                        Offsets p = lib[a].pts;
                        for (auto &q : lib[b].pts)
                            p.emplace_back(q.first + du, q.second + dv);
                        std::sort(p.begin(), p.end());
                        p.erase(std::unique(p.begin(), p.end()), p.end());
                        if (p.size() <= std::max(lib[a].pts.size(), lib[b].pts.size()))
                            continue;
                        // recipe relative to brick a's own origin: T_k[i] = max(T_a[i], T_b[i + (du,dv)])
                        const int ox = p[0].first, oy = p[0].second;
                        Offsets n = normalise(p);
                        bool dup = false;
                        for (auto &l : lib)
                            dup |= l.pts == n;
                        if (dup || cands.count(n))
                            continue;
                        (void)ox;
                        (void)oy;
                        cands[n] = {n, (int)a, (int)b, du, dv};
                    }
        std::vector<Brick> cl;
        for (auto &kv : cands)
            cl.push_back(kv.second);
        std::vector<long> score(cl.size(), 1L << 40);
#pragma omp parallel for schedule(dynamic, 8)
        for (size_t i = 0; i < cl.size(); ++i)
        {
            long tot = 0;
            bool any = false;
            for (size_t s = 0; s < sets.size(); ++s)
            {
                std::vector<Mask> c = cur[s];
                const size_t before = c.size();
                placements(sets[s], cl[i].pts, c);
                any |= c.size() > before;
                tot += greedy(c, (int)sets[s].S.size());
            }
            if (any)
                score[i] = tot;
        }
        size_t best = std::min_element(score.begin(), score.end()) - score.begin();
        lib.push_back(cl[best]);
        for (size_t s = 0; s < sets.size(); ++s)
            placements(sets[s], cl[best].pts, cur[s]);
        printf("K=%zu total %ld (%.2f per set)  brick %zu pts = a%d b%d (%d,%d):", lib.size(), score[best], (double)score[best] / sets.size(),
               cl[best].pts.size(), cl[best].a, cl[best].b, cl[best].du, cl[best].dv);
        for (auto &q : cl[best].pts)
            printf(" (%d,%d)", q.first, q.second);
        printf("\n");
        fflush(stdout);
    }
    return 0;
}
