// pack.cc — compacts the builder's tables into the device form: one shared-memory blob of class tables and the
// byte-offset entry descriptors (program.h).
#include <algorithm>
#include <array>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <stdexcept>

#include "structure.h"

namespace tb {

namespace {
int Align16(int v) { return (v + 15) & ~15; }
}

void PackProgram(HostProgram& hp) {
  // distinct local times and durations
  std::vector<double> times, durs;
  std::map<uint64_t, int> time_ix, dur_ix;
  auto intern = [](std::map<uint64_t, int>& ix, std::vector<double>& tab, double a, double b, double c) {
    uint64_t key;
    std::memcpy(&key, &a, 8);
    auto it = ix.find(key);
    if (it != ix.end()) return it->second;
    int i = static_cast<int>(tab.size() / 3);
    tab.push_back(a); tab.push_back(b); tab.push_back(c);
    ix[key] = i;
    return i;
  };
  std::vector<uint32_t> polys;
  for (const PolyRec& p : hp.polys) {
    int d = intern(dur_ix, durs, p.T, p.T2, p.T3);
    if (p.node0 > 0xFFFF || d > 0xFFFF) throw std::runtime_error("unsupported: spline too large for the compact tables");
    polys.push_back(static_cast<uint32_t>(p.node0) | (static_cast<uint32_t>(d) << 16));
  }
  std::vector<uint32_t> samples;
  auto add_sample = [&](int wide) {
    const SampleRec& r = hp.samples.at(wide);
    int t = intern(time_ix, times, r.t, r.t2, r.t3);
    // runtime samples of a PhaseSpline (optimised durations): poly = -(1 + RtSpline index), t = global time
    const int lo = r.poly < 0 ? -r.poly - 1 : r.poly;
    if (lo > 0xFFFF || t > 0xFFFF) throw std::runtime_error("unsupported: too many polynomials or sample times");
    samples.push_back(static_cast<uint32_t>(lo) | (static_cast<uint32_t>(t) << 16));
  };
  std::vector<ItemC> items;
  std::vector<std::array<int, 3>> kind;   // (level, type, sub) per compact item
  for (const ItemRec& it : hp.items) {
    ItemC c;
    std::memset(&c, 0, sizeof(c));
    c.type = static_cast<uint8_t>(it.type);
    c.sub = static_cast<uint8_t>(it.sub);
    c.ee = static_cast<uint8_t>(it.ee);
    c.n_ee = static_cast<uint8_t>(it.n_ee);
    if (it.gslot > 0xFFFF || it.b0 > 0xFFFF || it.aux > 0xFFFF) throw std::runtime_error("item slot out of range");
    c.gslot = static_cast<uint16_t>(it.gslot);
    c.b0 = static_cast<uint16_t>(it.b0);
    c.aux = static_cast<uint16_t>(it.aux);
    c.row = static_cast<uint32_t>(it.row[0]);
    size_t s0 = samples.size();
    switch (it.type) {
      case kItemDynPre: add_sample(it.s[kSBaseAng]); break;
      case kItemDynFoot:
        add_sample(it.s[kSBaseLin]);
        add_sample(it.s[kSMotion + it.ee]);
        add_sample(it.s[kSForce + it.ee]);
        add_sample(it.s[kSTorque + it.ee]);
        break;
      case kItemDynWrench: add_sample(it.s[kSBaseLin]); break;
      case kItemDynCol:
      case kItemDynJlin: s0 = 0; break;
      case kItemRomPre:
      case kItemBaseRom:
        add_sample(it.s[kSBaseLin]);
        add_sample(it.s[kSBaseAng]);
        break;
      case kItemTorque:
        add_sample(it.s[kSMotion + it.ee]);
        add_sample(it.s[kSForce + it.ee]);
        add_sample(it.s[kSTorque + it.ee]);
        break;
      case kItemRomFoot: add_sample(it.s[kSMotion + it.ee]); break;
      case kItemForce:
        add_sample(it.s[kSMotion + it.ee]);
        add_sample(it.s[kSForce + it.ee]);
        break;
      default:
        s0 = static_cast<size_t>(it.s[0]);   // node slot or polynomial index
        break;
    }
    if (s0 > 0xFFFF) throw std::runtime_error("unsupported: too many samples for the compact tables");
    c.s0 = static_cast<uint16_t>(s0);
    items.push_back(c);
    kind.push_back({ItemLevel(it.type), it.type, it.sub});
  }
  // per segment: level-0 items first, then level-1; items of one (type, sub) kind are contiguous and every kind
  // starts on a warp boundary of the two-lane kernel (16 items x 2 problem lanes) so that no warp of stage A diverges
  // between kinds (null items fill the gaps; one-lane kernels share a warp between two kinds at most)
  {
    std::vector<ItemC> padded;
    ItemC none;
    std::memset(&none, 0, sizeof(none));
    none.type = kItemNone;
    for (SegmentRec& sg : hp.segments) {
      const int new0 = static_cast<int>(padded.size());
      for (int level = 0; level < 2; ++level) {
        std::vector<std::array<int, 3>> kinds;
        for (int i = 0; i < sg.n_items; ++i) {
          auto k = kind[sg.item0 + i];
          if (k[0] != level) continue;
          bool seen = false;
          for (auto& q : kinds) seen |= (q == k);
          if (!seen) kinds.push_back(k);
        }
        // longest kinds first: where a level needs a second round of threads, that round holds the short items
        std::stable_sort(kinds.begin(), kinds.end(), [](const std::array<int, 3>& a, const std::array<int, 3>& b) {
          return ItemCostRank(a[1]) < ItemCostRank(b[1]);
        });
        for (auto& k : kinds) {
          while (padded.size() % kItemPad) padded.push_back(none);
          for (int i = 0; i < sg.n_items; ++i)
            if (kind[sg.item0 + i] == k) padded.push_back(items[sg.item0 + i]);
        }
        while (padded.size() % kItemPad) padded.push_back(none);
        if (level == 0) sg.n_items0 = static_cast<int>(padded.size()) - new0;
      }
      sg.item0 = new0;
      sg.n_items = static_cast<int>(padded.size()) - new0;
      if (std::getenv("TOWR_B200_DEBUG")) {
        std::fprintf(stderr, "segment %d kinds (type.sub x count):", static_cast<int>(&sg - hp.segments.data()));
        int run = 0, last_t = -1, last_s = -1, pos = new0;
        auto flush = [&]() { if (run) std::fprintf(stderr, " %d.%d x%d", last_t, last_s, run); };
        for (; pos < static_cast<int>(padded.size()); ++pos) {
          if (pos == new0 + sg.n_items0) { flush(); run = 0; std::fprintf(stderr, " |"); }
          const ItemC& c = padded[pos];
          if (c.type != last_t || c.sub != last_s) { flush(); run = 0; last_t = c.type; last_s = c.sub; }
          ++run;
        }
        flush();
        std::fprintf(stderr, "\n");
      }
    }
    items.swap(padded);
  }
  if (hp.n_vars >= 0xFFFF) throw std::runtime_error("unsupported: more than 65534 variables");

  BlobLayout& L = hp.blob;
  int off = 0;
  auto place = [&](int& o, int bytes) { o = off; off = Align16(off + bytes); };
  L.n_B = static_cast<int>(hp.B.size());
  L.n_times = static_cast<int>(times.size() / 3);
  L.n_durs = static_cast<int>(durs.size() / 3);
  L.n_items = static_cast<int>(items.size());
  L.n_samples = static_cast<int>(samples.size());
  L.n_polys = static_cast<int>(polys.size());
  L.n_node_values = hp.n_node_values;
  place(L.off_B, 8 * L.n_B);
  place(L.off_times, 24 * L.n_times);
  place(L.off_durs, 24 * L.n_durs);
  place(L.off_rdurs, 16 * L.n_durs);
  L.n_segments = static_cast<int>(hp.segments.size());
  place(L.off_segments, static_cast<int>(sizeof(SegmentRec)) * L.n_segments);
  place(L.off_items, 16 * L.n_items);
  place(L.off_samples, 4 * L.n_samples);
  place(L.off_polys, 4 * L.n_polys);
  place(L.off_node_src, 2 * L.n_node_values);
  L.rt = hp.optimize_durations ? 1 : 0;
  L.n_rt_splines = static_cast<int>(hp.rt_splines.size());
  L.n_rt_feet = static_cast<int>(hp.rt_feet.size());
  L.n_pdur = 0;
  for (const RtFoot& f : hp.rt_feet) L.n_pdur += f.n_phases;
  place(L.off_rt_splines, static_cast<int>(sizeof(RtSpline)) * L.n_rt_splines);
  place(L.off_rt_feet, static_cast<int>(sizeof(RtFoot)) * L.n_rt_feet);
  place(L.off_poly_meta, 4 * L.n_polys);
  L.bytes = off;
  hp.blob_data.assign(off, 0);
  unsigned char* d = hp.blob_data.data();
  std::memcpy(d + L.off_B, hp.B.data(), 8 * hp.B.size());
  std::memcpy(d + L.off_times, times.data(), 8 * times.size());
  std::memcpy(d + L.off_durs, durs.data(), 8 * durs.size());
  {
    double* rd = reinterpret_cast<double*>(d + L.off_rdurs);
    for (int i = 0; i < L.n_durs; ++i) { rd[2 * i] = 1.0 / durs[3 * i + 1]; rd[2 * i + 1] = 1.0 / durs[3 * i + 2]; }
  }
  std::memcpy(d + L.off_segments, hp.segments.data(), sizeof(SegmentRec) * hp.segments.size());
  std::memcpy(d + L.off_items, items.data(), 16 * items.size());
  std::memcpy(d + L.off_samples, samples.data(), 4 * samples.size());
  std::memcpy(d + L.off_polys, polys.data(), 4 * polys.size());
  if (!hp.rt_splines.empty()) std::memcpy(d + L.off_rt_splines, hp.rt_splines.data(), sizeof(RtSpline) * hp.rt_splines.size());
  if (!hp.rt_feet.empty()) std::memcpy(d + L.off_rt_feet, hp.rt_feet.data(), sizeof(RtFoot) * hp.rt_feet.size());
  std::memcpy(d + L.off_poly_meta, hp.poly_meta.data(), 4 * hp.poly_meta.size());
  uint16_t* ns = reinterpret_cast<uint16_t*>(d + L.off_node_src);
  for (int i = 0; i < hp.n_node_values; ++i) ns[i] = hp.node_src[i] < 0 ? 0xFFFF : static_cast<uint16_t>(hp.node_src[i]);

  hp.dyn_index.clear();
  hp.dyn_index_flat.clear();
  for (size_t e = 0; e < hp.entries.size(); ++e) {
    const uint32_t w = hp.entries[e];
    if ((w & kEntryGG) || static_cast<int>(w & 0xFFFF) >= kGReserved) {
      hp.dyn_index.push_back(static_cast<int>(e));
      if (hp.entry_code[e] < 0) hp.dyn_index_flat.push_back(static_cast<int>(e));
    }
  }
  if (std::getenv("TOWR_B200_DEBUG")) {
    auto runs = [](const std::vector<int>& v) { int r = v.empty() ? 0 : 1; for (size_t i = 1; i < v.size(); ++i) r += v[i] != v[i - 1] + 1; return r; };
    std::fprintf(stderr, "dynamic entries: %zu of %zu in %d runs; on zero-slope terrains %zu in %d runs\n", hp.dyn_index.size(),
                 hp.entries.size(), runs(hp.dyn_index), hp.dyn_index_flat.size(), runs(hp.dyn_index_flat));
    if (std::atoi(std::getenv("TOWR_B200_DEBUG")) > 1) {
      const std::vector<int>& v = hp.dyn_index_flat;
      std::fprintf(stderr, "flat runs (pos:len):");
      for (size_t k = 0; k < v.size();) { size_t e = k + 1; while (e < v.size() && v[e] == v[e - 1] + 1) ++e; std::fprintf(stderr, " %d:%zu", v[k], e - k); k = e; }
      std::fprintf(stderr, "\n");
    }
  }
  // device descriptors: byte offsets, so that stage B needs no index scaling
  hp.entries_dev.clear();
  hp.entries_dev.reserve(hp.entries.size());
  for (uint32_t w : hp.entries) {
    uint32_t g8 = (w & 0xFFFF) * 8, b8 = ((w >> 16) & 0x7FFF) * 8;
    if (g8 > 0xFFFF || b8 > 0xFFFF) throw std::runtime_error("unsupported: G or B table beyond 8191 entries");
    hp.entries_dev.push_back(g8 | ((b8 | ((w & kEntryGG) ? 1u : 0u)) << 16));
  }
}

}  // namespace tb
