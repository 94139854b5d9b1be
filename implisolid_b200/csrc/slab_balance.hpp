// z-slab boundaries: integer arithmetic shared by every rank (geometry_api.cu: balanced_layers) — host only, no CUDA, so that the
// CPU suite runs the very code the library runs (tests/emu, tests/test_slab_balance.py).
//
// Weights: one per cube layer.  First the count of surface tiles in the layer's bin of eight node planes plus a uniform term
// (a quarter of the mean, the per-layer fixed costs), as whole numbers << 16; cuts at equal cumulative weight, at least one layer
// per rank (implisolid_b200/slabs.py: balanced_cuts mirrors this first cut).  Feedback: given every rank's measured busy time
// under the current cuts, each layer's weight is multiplied by its owner's cost per unit of weight relative to the mean
// (x / 1024 fixed point, clamped to [1/4, 4]) and the cuts are drawn again.  Everything is integer, so ranks that hold the same
// histogram and the same table of times draw the same cuts without talking to each other.
#pragma once
#include <cstdint>
#include <vector>

namespace isd {

struct SlabBalance {
    std::vector<unsigned long long> weight;      // per cube layer
    std::vector<int> cuts;                       // world + 1 entries: rank r owns cube layers [cuts[r], cuts[r + 1])
    int gen = 0, recuts = 0;
    bool scaled = false;                         // feedback factors have been multiplied in

    void draw(int world) {
        const int layers = (int)weight.size(), shift = scaled ? 0 : 16;
        std::vector<unsigned long long> cum((size_t)layers + 1, 0ull);
        for (int l = 0; l < layers; ++l) cum[(size_t)l + 1] = cum[(size_t)l] + (weight[(size_t)l] >> shift);
        cuts.assign((size_t)world + 1, 0);
        cuts[(size_t)world] = layers;
        for (int r = 1; r < world; ++r) {
            const unsigned long long target = cum[(size_t)layers] * (unsigned long long)r / (unsigned long long)world;
            int l = cuts[(size_t)r - 1] + 1;                                   // at least one layer per rank
            while (l < layers - (world - r) && cum[(size_t)l] < target) ++l;
            cuts[(size_t)r] = l;
        }
    }
    // hist[b] = surface tiles whose node planes are [8b, 8b + 8); cube layer l (0-based, node plane l + 1) falls in bin (l + 1) >> 3
    void init(const unsigned* hist, int bins, int layers, int world) {
        unsigned long long total_bins = 0;
        for (int b = 0; b < bins; ++b) total_bins += hist[b];
        const unsigned long long uniform = total_bins * 8ull / (4ull * (unsigned long long)layers) + 1ull;
        weight.assign((size_t)layers, 0ull);
        for (int l = 0; l < layers; ++l) weight[(size_t)l] = ((unsigned long long)hist[(size_t)((l + 1) >> 3)] + uniform) << 16;
        gen = 0; recuts = 0; scaled = false;
        draw(world);
    }
    // busy[r] > 0: rank r's time under the current cuts.  Returns true (and bumps gen) if the cuts were drawn again.
    bool feedback(const long long* busy, int world, int threshold_permille) {
        const int layers = (int)weight.size();
        long long hi = 0, sum = 0;
        for (int r = 0; r < world; ++r) { if (busy[r] <= 0) return false; hi = busy[r] > hi ? busy[r] : hi; sum += busy[r]; }
        if ((__int128)hi * world * 1000 <= (__int128)sum * threshold_permille) return false;
        unsigned long long wsum = 0;
        for (int l = 0; l < layers; ++l) wsum += weight[(size_t)l] >> 16;
        for (int r = 0; r < world; ++r) {
            unsigned long long w = 0;
            for (int l = cuts[(size_t)r]; l < cuts[(size_t)r + 1]; ++l) w += weight[(size_t)l] >> 16;
            if (w == 0) continue;
            // (t_r / w_r) / (sum t / sum w) x 1024
            unsigned long long f = (unsigned long long)(((__int128)busy[r] * (__int128)wsum * 1024) / ((__int128)sum * (__int128)w));
            f = f < 256 ? 256 : (f > 4096 ? 4096 : f);      // one noisy step must not fling the cuts away
            for (int l = cuts[(size_t)r]; l < cuts[(size_t)r + 1]; ++l) {
                const unsigned long long nw = (weight[(size_t)l] >> 10) * f;
                weight[(size_t)l] = nw > 0 ? nw : 1ull;
            }
        }
        // keep the weights in range over many re-cuts: halve everything while the largest is above 2^44
        unsigned long long mx = 0;
        for (unsigned long long v : weight) mx = v > mx ? v : mx;
        while (mx >= (1ull << 44)) { for (unsigned long long& v : weight) v = v > 1 ? v >> 1 : 1ull; mx >>= 1; }
        scaled = true;
        draw(world);
        ++gen; ++recuts;
        return true;
    }
};

}  // namespace isd
