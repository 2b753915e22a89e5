// ORACLE — TEST INFRASTRUCTURE ONLY (see towr_oracle.h header).
// Gait generators: named gaits -> per-foot phase durations and initial contact flags.
#include <numeric>
#include <stdexcept>

#include "towr_oracle.h"

namespace orc {

// ============================================================ gait_generator.cc
// :54-64
Vec GaitGenerator::GetPhaseDurations(double t_total, int ee) const {
  auto v = GetPhaseDurations().at(ee);               // :66-75 normalised, then scaled
  double total_time = std::accumulate(v.begin(), v.end(), 0.0);
  for (double& t_phase : v) t_phase = t_phase / total_time;
  Vec durations;
  for (double d : v) durations.push_back(d * t_total);
  return durations;
}
// :77-105
std::vector<Vec> GaitGenerator::GetPhaseDurations() const {
  int n_ee = static_cast<int>(contacts_.front().size());
  Vec d_accumulated(n_ee, 0.0);
  std::vector<Vec> foot_durations(n_ee);
  for (size_t phase = 0; phase + 1 < contacts_.size(); ++phase) {
    const ContactState& curr = contacts_.at(phase);
    const ContactState& next = contacts_.at(phase + 1);
    for (size_t ee = 0; ee < curr.size(); ++ee) {
      d_accumulated.at(ee) += times_.at(phase);
      bool contacts_will_change = curr.at(ee) != next.at(ee);
      if (contacts_will_change) {
        foot_durations.at(ee).push_back(d_accumulated.at(ee));
        d_accumulated.at(ee) = 0.0;
      }
    }
  }
  for (size_t ee = 0; ee < contacts_.back().size(); ++ee)
    foot_durations.at(ee).push_back(d_accumulated.at(ee) + times_.back());
  return foot_durations;
}
// :113-129
void GaitGenerator::SetGaits(const std::vector<Gaits>& gaits) {
  contacts_.clear();
  times_.clear();
  for (Gaits g : gaits) {
    auto info = GetGait(g);
    times_.insert(times_.end(), info.first.begin(), info.first.end());
    contacts_.insert(contacts_.end(), info.second.begin(), info.second.end());
  }
}
// :131-144
GaitGenerator::GaitInfo GaitGenerator::RemoveTransition(const GaitInfo& g) const {
  GaitInfo new_gait = g;
  new_gait.first.pop_back();
  new_gait.first.back() += g.first.back();
  new_gait.second.pop_back();
  return new_gait;
}

namespace {
using CS = GaitGenerator::ContactState;

// monoped_gait_generator.cc:37-120
struct MonopedGaitGenerator : GaitGenerator {
  CS o_{true}, x_{false};
  MonopedGaitGenerator() { SetGaits({Stand}); }
  void SetCombo(Combos combo) override {
    switch (combo) {
      case C0: SetGaits({Stand, Hop1, Hop1, Hop1, Hop1, Stand}); break;
      case C1: SetGaits({Stand, Hop1, Hop1, Hop1, Stand}); break;
      case C2: SetGaits({Stand, Hop1, Hop1, Hop1, Hop1, Stand}); break;
      case C3: SetGaits({Stand, Hop2, Hop2, Hop2, Stand}); break;
      case C4: SetGaits({Stand, Hop2, Hop2, Hop2, Hop2, Hop2, Stand}); break;
    }
  }
  GaitInfo GetGait(Gaits gait) const override {
    switch (gait) {
      case Stand: return {{0.5}, {o_}};
      case Flight: return {{0.5}, {x_}};
      case Hop1: return {{0.3, 0.3}, {o_, x_}};
      case Hop2: return {{0.2, 0.3}, {o_, x_}};
      default: throw std::runtime_error("gait not implemented");
    }
  }
};

// biped_gait_generator.cc:38-226 ; ids L=0, R=1
struct BipedGaitGenerator : GaitGenerator {
  CS I_{false, false}, b_{false, true}, P_{true, false}, B_{true, true};
  BipedGaitGenerator() { SetGaits({Stand}); }
  void SetCombo(Combos combo) override {
    switch (combo) {
      case C0: SetGaits({Stand, Walk1, Walk1, Walk1, Walk1, Stand}); break;
      case C1: SetGaits({Stand, Run1, Run1, Run1, Run1, Stand}); break;
      case C2: SetGaits({Stand, Hop1, Hop1, Hop1, Stand}); break;
      case C3: SetGaits({Stand, Hop1, Hop2, Hop2, Stand}); break;
      case C4: SetGaits({Stand, Hop5, Hop5, Hop5, Stand}); break;
    }
  }
  GaitInfo GetGait(Gaits gait) const override {
    switch (gait) {
      case Stand: return {{0.2}, {B_}};
      case Flight: return {{0.5}, {I_}};
      case Walk1: case Walk2: { double step = 0.3, stance = 0.05; return {{step, stance, step, stance}, {b_, B_, P_, B_}}; }
      case Run1: case Run3: { double flight = 0.4, pushoff = 0.15, landing = 0.15;
        return {{pushoff, flight, landing + pushoff, flight, landing}, {b_, I_, P_, I_, b_}}; }
      case Hop1: return {{0.15, 0.5, 0.15}, {B_, I_, B_}};
      case Hop2: return {{0.15, 0.4, 0.15}, {b_, I_, b_}};
      case Hop3: return {{0.2, 0.2, 0.2}, {P_, I_, P_}};
      case Hop5: { double push = 0.2, flight = 0.3, land = 0.2; return {{push, flight, land, land}, {P_, I_, b_, B_}}; }
      default: throw std::runtime_error("gait not implemented");
    }
  }
};

// quadruped_gait_generator.cc:38-366 ; ids LF=0, RF=1, LH=2, RH=3
struct QuadrupedGaitGenerator : GaitGenerator {
  CS II_, PI_, bI_, IP_, Ib_, Pb_, bP_, BI_, IB_, PP_, bb_, Bb_, BP_, bB_, PB_, BB_;
  static CS mk(std::initializer_list<int> on) { CS c(4, false); for (int i : on) c[i] = true; return c; }
  QuadrupedGaitGenerator() {
    const int LF = 0, RF = 1, LH = 2, RH = 3;
    II_ = mk({});
    PI_ = mk({LH}); bI_ = mk({RH}); IP_ = mk({LF}); Ib_ = mk({RF});
    Pb_ = mk({LH, RF}); bP_ = mk({RH, LF}); BI_ = mk({LH, RH}); IB_ = mk({LF, RF}); PP_ = mk({LH, LF}); bb_ = mk({RH, RF});
    Bb_ = mk({LH, RH, RF}); BP_ = mk({LH, RH, LF}); bB_ = mk({RH, LF, RF}); PB_ = mk({LH, LF, RF});
    BB_ = mk({LF, RF, LH, RH});
    SetGaits({Stand});
  }
  void SetCombo(Combos combo) override {
    switch (combo) {
      case C0: SetGaits({Stand, Walk2, Walk2, Walk2, Walk2E, Stand}); break;
      case C1: SetGaits({Stand, Run2, Run2, Run2, Run2E, Stand}); break;
      case C2: SetGaits({Stand, Run3, Run3, Run3, Run3E, Stand}); break;
      case C3: SetGaits({Stand, Hop1, Hop1, Hop1, Hop1E, Stand}); break;
      case C4: SetGaits({Stand, Hop3, Hop3, Hop3, Hop3E, Stand}); break;
    }
  }
  GaitInfo WalkOverlap() const { double three = 0.25, lateral = 0.13, diagonal = 0.13;
    return {{three, lateral, three, diagonal, three, lateral, three, diagonal}, {bB_, bb_, Bb_, Pb_, PB_, PP_, BP_, bP_}}; }
  GaitInfo Gallop() const { double A = 0.3, B = 0.2, C = 0.2;
    return {{B, A, B, C, B, A, B, C}, {Bb_, BI_, BP_, bP_, bB_, IB_, PB_, Pb_}}; }
  GaitInfo GetGait(Gaits gait) const override {
    switch (gait) {
      case Stand: return {{0.3}, {BB_}};
      case Flight: return {{0.3}, {Bb_}};                                  // :126-140 (sic: Bb_)
      case Walk1: { double step = 0.3, stand = 0.2;
        return {{step, stand, step, stand, step, stand, step, stand}, {bB_, BB_, Bb_, BB_, PB_, BB_, BP_, BB_}}; }
      case Walk2: return WalkOverlap();
      case Walk2E: return RemoveTransition(WalkOverlap());
      case Run1: { double t_step = 0.3, t_stand = 0.2; return {{t_step, t_stand, t_step, t_stand}, {bP_, BB_, Pb_, BB_}}; }
      case Run2: { double stand = 0.4, flight = 0.1; return {{stand, flight, stand, flight}, {bP_, II_, Pb_, II_}}; }
      case Run2E: return {{0.4}, {bP_}};
      case Run3: { double stand = 0.3, flight = 0.1; return {{stand, flight, stand, flight}, {PP_, II_, bb_, II_}}; }
      case Run3E: return {{0.3}, {PP_}};
      case Hop1: { double stand = 0.3, flight = 0.1; return {{stand, flight, stand, flight}, {BI_, II_, IB_, II_}}; }
      case Hop1E: return {{0.3}, {BI_}};
      case Hop2: return {{0.3, 0.4, 0.3}, {BB_, II_, BB_}};
      case Hop3: return Gallop();
      case Hop3E: return RemoveTransition(Gallop());
      case Hop5: { double A = 0.1, B = 0.2, C = 0.1; return {{A, B, C, A, B, C}, {Bb_, BB_, IP_, Bb_, BB_, IP_}}; }
      default: throw std::runtime_error("gait not implemented");
    }
  }
};
}  // namespace

// gait_generator.cc:43-52
std::shared_ptr<GaitGenerator> GaitGenerator::MakeGaitGenerator(int leg_count) {
  switch (leg_count) {
    case 1: return std::make_shared<MonopedGaitGenerator>();
    case 2: return std::make_shared<BipedGaitGenerator>();
    case 4: return std::make_shared<QuadrupedGaitGenerator>();
    default: throw std::runtime_error("gait generator not implemented for this leg count");
  }
}

}  // namespace orc
