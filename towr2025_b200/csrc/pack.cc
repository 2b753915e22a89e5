// pack.cc — compacts the builder's tables into the device form: shared-memory blobs of class tables and the
// byte-offset entry descriptors (program.h).
//
// A class whose tables fit one SM is ONE part (one blob, one kernel launch per evaluation).  A class with many samples
// (sampling steps of 0.02 s and below) is split into PARTS: runs of consecutive segments, each with its own blob (its
// items, samples, sample times and B constants; the spline tables every part needs are repeated) and its own launch.
// Parts write disjoint g rows and disjoint ranges of the value array; stages N / C are recomputed per part.
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

// B constants an item reads by position from ItemRec::b0 (structure.cc): the dynamic sample's 16 Hermite basis values
// (every item derived from the sample carries the same b0), the torque constraint's k_friction
int ItemBRun(int type, int sub) {
  switch (type) {
    case kItemTerrainHard: return 3 * sub;               // (J_pos, J_vel, dim) per stored column; 0 with optimised durations
    case kItemDynPre: case kItemDynFoot: case kItemDynWrench: case kItemDynJlin: case kItemDynCol: return 16;
    case kItemTorque: return 1;
    case kItemEELinear: return 4 * sub;                  // (coeff, dimension, derivative, G offset of the foot's block) per term
    default: return 0;
  }
}

struct Unfit : std::runtime_error {   // this range of segments does not fit the compact tables: try a shorter range
  using std::runtime_error::runtime_error;
};

// packs segments [seg0, seg1) of `wide` (the builder's unpadded segment list) into `part` and writes the descriptors of
// their entries into hp.entries_dev
void PackRange(HostProgram& hp, const std::vector<SegmentRec>& wide, int seg0, int seg1, ProgramPart& part) {
  // distinct local times and durations
  std::vector<double> times, durs;
  // records are shared by value of ALL three doubles: a run-time sample (global time, 0, 0) and a fixed-duration sample
  // whose local time has the same bits (local time, t^2, t^3) are different records
  using Key3 = std::array<uint64_t, 3>;
  std::map<Key3, int> time_ix, dur_ix;
  auto intern = [](std::map<Key3, int>& ix, std::vector<double>& tab, double a, double b, double c) {
    Key3 key;
    std::memcpy(&key[0], &a, 8);
    std::memcpy(&key[1], &b, 8);
    std::memcpy(&key[2], &c, 8);
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
  // B of this part: the constants its entries multiply by first (their byte offsets must fit 16 bits), then the runs
  // its items read by position
  std::vector<double> Bp;
  std::map<int, int> b_local;                            // global B index -> local
  const int e_begin = wide[seg0].entry0, e_end = wide[seg1 - 1].entry0 + wide[seg1 - 1].n_entries;
  for (int e = e_begin; e < e_end; ++e) {
    const uint32_t w = hp.entries[e];
    if (w & kEntryGG) continue;
    const int b = static_cast<int>((w >> 16) & 0x7FFF);
    if (b_local.emplace(b, static_cast<int>(Bp.size())).second) Bp.push_back(hp.B.at(b));
  }
  if (Bp.size() > 8191) throw Unfit("B table beyond 8191 entries");
  std::map<int, int> run_local;                          // global b0 -> local b0
  auto local_run = [&](int b0, int len) {
    auto it = run_local.find(b0);
    if (it != run_local.end()) return it->second;
    const int at = static_cast<int>(Bp.size());
    for (int i = 0; i < len; ++i) Bp.push_back(hp.B.at(b0 + i));
    run_local[b0] = at;
    return at;
  };

  std::vector<uint32_t> samples;
  auto add_sample = [&](int wide_ix) {
    const SampleRec& r = hp.samples.at(wide_ix);
    int t = intern(time_ix, times, r.t, r.t2, r.t3);
    // runtime samples of a PhaseSpline (optimised durations): poly = -(1 + RtSpline index), t = global time
    const int lo = r.poly < 0 ? -r.poly - 1 : r.poly;
    if (lo > 0xFFFF) throw std::runtime_error("unsupported: too many polynomials for the compact tables");
    if (t > 0xFFFF) throw Unfit("too many sample times");
    samples.push_back(static_cast<uint32_t>(lo) | (static_cast<uint32_t>(t) << 16));
  };
  const int item_begin = wide[seg0].item0, item_end = wide[seg1 - 1].item0 + wide[seg1 - 1].n_items;
  std::vector<ItemC> items;                              // compact items of [item_begin, item_end)
  std::vector<std::array<int, 3>> kind;                  // (level, type, sub) per compact item
  for (int ii = item_begin; ii < item_end; ++ii) {
    const ItemRec& it = hp.items[ii];
    ItemC c;
    std::memset(&c, 0, sizeof(c));
    c.type = static_cast<uint8_t>(it.type);
    c.sub = static_cast<uint8_t>(it.sub);
    c.ee = static_cast<uint8_t>(it.ee);
    c.n_ee = static_cast<uint8_t>(it.n_ee);
    const int run = ItemBRun(it.type, it.sub);
    const int b0 = run ? local_run(it.b0, run) : (it.type == kItemTorqueNode ? it.b0 : 0);   // TorqueNode: a node slot
    if (it.gslot > 0xFFFF || it.aux > 0xFFFF) throw std::runtime_error("item slot out of range");
    if (b0 > 0xFFFF) throw Unfit("B runs beyond 65535 entries");
    c.gslot = static_cast<uint16_t>(it.gslot);
    c.b0 = static_cast<uint16_t>(b0);
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
      case kItemTerrainHard: add_sample(it.s[kSMotion + it.ee]); break;
      case kItemEELinear:
        for (int j = 0; j < it.sub; ++j) add_sample(it.s[j]);   // one sample per term, in term order
        break;
      case kItemForce:
        add_sample(it.s[kSMotion + it.ee]);
        add_sample(it.s[kSForce + it.ee]);
        break;
      default:
        s0 = static_cast<size_t>(it.s[0]);   // node slot or polynomial index
        break;
    }
    if (s0 > 0xFFFF) throw Unfit("too many samples for the compact tables");
    c.s0 = static_cast<uint16_t>(s0);
    items.push_back(c);
    kind.push_back({ItemLevel(it.type), it.type, it.sub});
  }
  // per segment: level-0 items first, then level-1; items of one (type, sub) kind are contiguous and every kind
  // starts on a warp boundary of the two-lane kernel (16 items x 2 problem lanes) so that no warp of stage A diverges
  // between kinds (null items fill the gaps; one-lane kernels share a warp between two kinds at most)
  part.segments.assign(wide.begin() + seg0, wide.begin() + seg1);
  part.g_capacity = 0;
  {
    std::vector<ItemC> padded;
    ItemC none;
    std::memset(&none, 0, sizeof(none));
    none.type = kItemNone;
    for (SegmentRec& sg : part.segments) {
      part.g_capacity = std::max(part.g_capacity, sg.g_size);
      const int first = sg.item0 - item_begin;           // in `items`
      const int new0 = static_cast<int>(padded.size());
      for (int level = 0; level < 2; ++level) {
        std::vector<std::array<int, 3>> kinds;
        for (int i = 0; i < sg.n_items; ++i) {
          auto k = kind[first + i];
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
            if (kind[first + i] == k) padded.push_back(items[first + i]);
        }
        while (padded.size() % kItemPad) padded.push_back(none);
        if (level == 0) sg.n_items0 = static_cast<int>(padded.size()) - new0;
      }
      sg.item0 = new0;
      sg.n_items = static_cast<int>(padded.size()) - new0;
      if (std::getenv("TOWR_B200_DEBUG")) {
        std::fprintf(stderr, "segment %d kinds (type.sub x count):", seg0 + static_cast<int>(&sg - part.segments.data()));
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

  BlobLayout& L = part.blob;
  L = BlobLayout{};
  int off = 0;
  auto place = [&](int& o, int bytes) { o = off; off = Align16(off + bytes); };
  L.n_B = static_cast<int>(Bp.size());
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
  L.n_segments = static_cast<int>(part.segments.size());
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
  part.blob_data.assign(off, 0);
  unsigned char* d = part.blob_data.data();
  std::memcpy(d + L.off_B, Bp.data(), 8 * Bp.size());
  std::memcpy(d + L.off_times, times.data(), 8 * times.size());
  std::memcpy(d + L.off_durs, durs.data(), 8 * durs.size());
  {
    double* rd = reinterpret_cast<double*>(d + L.off_rdurs);
    for (int i = 0; i < L.n_durs; ++i) { rd[2 * i] = 1.0 / durs[3 * i + 1]; rd[2 * i + 1] = 1.0 / durs[3 * i + 2]; }
  }
  std::memcpy(d + L.off_segments, part.segments.data(), sizeof(SegmentRec) * part.segments.size());
  std::memcpy(d + L.off_items, items.data(), 16 * items.size());
  std::memcpy(d + L.off_samples, samples.data(), 4 * samples.size());
  std::memcpy(d + L.off_polys, polys.data(), 4 * polys.size());
  if (!hp.rt_splines.empty()) std::memcpy(d + L.off_rt_splines, hp.rt_splines.data(), sizeof(RtSpline) * hp.rt_splines.size());
  if (!hp.rt_feet.empty()) std::memcpy(d + L.off_rt_feet, hp.rt_feet.data(), sizeof(RtFoot) * hp.rt_feet.size());
  std::memcpy(d + L.off_poly_meta, hp.poly_meta.data(), 4 * hp.poly_meta.size());
  uint16_t* ns = reinterpret_cast<uint16_t*>(d + L.off_node_src);
  for (int i = 0; i < hp.n_node_values; ++i) ns[i] = hp.node_src[i] < 0 ? 0xFFFF : static_cast<uint16_t>(hp.node_src[i]);

  // device descriptors of this part's entries: byte offsets, so that stage B needs no index scaling
  for (int e = e_begin; e < e_end; ++e) {
    const uint32_t w = hp.entries[e];
    const bool gg = (w & kEntryGG) != 0;
    const uint32_t second = gg ? ((w >> 16) & 0x7FFF) : static_cast<uint32_t>(b_local.at(static_cast<int>((w >> 16) & 0x7FFF)));
    const uint32_t g8 = (w & 0xFFFF) * 8, b8 = second * 8;
    if (g8 > 0xFFFF || b8 > 0xFFFF) throw Unfit("G or B table beyond 8191 entries");
    hp.entries_dev[e] = g8 | ((b8 | (gg ? 1u : 0u)) << 16);
  }
}
}  // namespace

void PackProgram(HostProgram& hp) {
  if (hp.n_vars >= 0xFFFF) throw std::runtime_error("unsupported: more than 65534 variables");
  const std::vector<SegmentRec> wide = hp.segments;      // the builder's list: item0 / n_items index hp.items
  const int n_seg = static_cast<int>(wide.size());
  hp.entries_dev.assign(hp.entries.size(), kNoEntry);
  hp.parts.clear();
  // One part when the whole class fits the single-blob budget; otherwise greedy runs of segments under the per-part
  // budget (TOWR_B200_MAX_BLOB overrides both: tests use it to force several parts on small classes).
  int single_budget = 150 * 1024, part_budget = 72 * 1024;
  if (const char* env = std::getenv("TOWR_B200_MAX_BLOB")) single_budget = part_budget = std::max(4096, std::atoi(env));
  bool done = false;
  if (n_seg > 0) {
    try {
      ProgramPart all;
      PackRange(hp, wide, 0, n_seg, all);
      if (all.blob.bytes <= single_budget) {
        hp.parts.push_back(std::move(all));
        done = true;
      }
    } catch (const Unfit&) {
    }
  }
  for (int a = 0; !done && a < n_seg;) {
    ProgramPart best;
    int b = a;
    while (b < n_seg) {
      ProgramPart trial;
      try {
        PackRange(hp, wide, a, b + 1, trial);
      } catch (const Unfit& e) {
        if (b == a) throw std::runtime_error(std::string("unsupported: one segment alone exceeds the compact tables: ") + e.what());
        break;
      }
      if (b > a && trial.blob.bytes > part_budget) break;
      best = std::move(trial);
      ++b;
    }
    PackRange(hp, wide, a, b, best);                     // rewrite the descriptors a failed longer trial may have touched
    hp.parts.push_back(std::move(best));
    a = b;
  }
  if (hp.parts.empty()) hp.parts.emplace_back();
  // single-part view (host emulation, debug prints, sizing of the common case)
  hp.blob = hp.parts.front().blob;
  hp.blob_data = hp.parts.front().blob_data;
  hp.segments.clear();
  for (const ProgramPart& p : hp.parts) hp.segments.insert(hp.segments.end(), p.segments.begin(), p.segments.end());

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
    std::fprintf(stderr, "parts: %zu (blob bytes:", hp.parts.size());
    for (const ProgramPart& p : hp.parts) std::fprintf(stderr, " %d/%zu seg", p.blob.bytes, p.segments.size());
    std::fprintf(stderr, ")\n");
    if (std::atoi(std::getenv("TOWR_B200_DEBUG")) > 1) {
      const std::vector<int>& v = hp.dyn_index_flat;
      std::fprintf(stderr, "flat runs (pos:len):");
      for (size_t k = 0; k < v.size();) { size_t e = k + 1; while (e < v.size() && v[e] == v[e - 1] + 1) ++e; std::fprintf(stderr, " %d:%zu", v[k], e - k); k = e; }
      std::fprintf(stderr, "\n");
    }
  }
}

}  // namespace tb
