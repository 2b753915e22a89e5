// program.h — the "evaluation program" of one structure class.
//
// A structure class fixes everything the sparsity pattern of TOWR's constraint Jacobian depends on
// (robot leg count, phase counts, contact-at-start, fixed phase durations, sampling dt's, constraint
// list).  The host builder (structure.cc) turns a towr_b200_spec into flat tables; the kernel
// (kernels.cu) interprets them for every problem of a batch:
//
//   stage N  gather node values (pos/vel per spline node) from x            -> smem nodes[]
//   stage C  Hermite coefficients c,d per (polynomial, dim)                 -> smem coef[]
//   per segment:
//     stage A  one thread per *item* (a time sample or a node of one constraint family):
//              writes its constraint rows g and its partial-derivative slots -> smem G[]
//     stage B  one thread per stored Jacobian entry: val = G[g] * B[b], coalesced store into the problem's
//              value array (entries that are sums of products are finished by their item in stage A)
//
// B[] holds class constants (Hermite basis values at the fixed sample times, signs, 1/t_swing ...),
// G[] holds per-problem quantities (forces, lever arms, rotation matrices, terrain slopes ...).
#pragma once
#include <cstdint>
#include <cstddef>
#include <string>
#ifndef __CUDACC__
struct alignas(16) uint4 { unsigned int x, y, z, w; };
#endif
#include <vector>

#include "terrain.h"

namespace tb {

// ---- device-visible records (plain data, shared by host and device code) ---------------------------------------
struct PolyRec {        // one cubic-Hermite polynomial of one spline
  int node0;            // slot of its start node in nodes[] (end node = node0 + 1)
  int pad;
  double T, T2, T3;     // duration and std::pow(T,2), std::pow(T,3) (host-evaluated, polynomial.cc:97-104)
};

struct SampleRec {      // "spline S evaluated at global time t" for fixed durations (spline.cc:48-78)
  int poly;             // global polynomial index
  int pad;
  double t, t2, t3;     // local time and std::pow(t,2), std::pow(t,3)
};

enum ItemType : int {
  // level 0: independent of other items
  kItemDynPre = 0,      // per dynamic sample: euler point, trig, R, R I_b, I_w, M, Mdot, omega, omegadot ... -> scratch slots
  kItemRomPre,          // per RoM sample: base position, R and dR/d(euler) -> scratch slots
  kItemForce,           // force_constraint_discretized.cc:97-221  5 rows per (sample, foot)
  kItemTerrainNode,     // terrain_constraint.cc:61-111            1 row per ee-motion node
  kItemBaseHeightNode,  // base_height_constraint.cc:58-110        1 row per base-lin node
  kItemSwingNode,       // swing_constraint.cc:54-108              4 rows per swing node
  kItemSplineAcc,       // spline_acc_constraint.cc:48-80          3 rows per junction
  kItemTotalDuration,   // total_duration_constraint.cc:48-72
  kItemForceNode,       // force_constraint.cc:62-171                5 rows per stance force node (node-based variant)
  kItemBaseRom,         // base_motion_constraint.cc:60-85         6 rows per sample (base euler angles, base position)
  kItemTorque,          // torque_constraint_discretized.cc:105-235  4 rows per (sample, foot)
  kItemDynFoot,         // per (dynamic sample, foot): f_e, c - p_e slots and the foot's torque term f_e x (c - p_e) + tau_e
  kItemTerrainHard,     // terrain_constraint_hard.cc:50-132         1 row per (sample, foot), finished ee-motion columns
  kItemTorqueNode,      // torque_constraint.cc:64-157               3 rows per stance torque node (node-based variant)
  kItemEELinear,        // ee_linear_constraint.cc:20-29             1 row per sample: sum of coeff * (foot spline value)
  // level 1: read the scratch slots of a level-0 item (ItemC::aux)
  kItemDynWrench,       // dynamic_constraint.cc:63-148: the 6 g rows of a dynamic sample (feet summed in order)
  kItemDynJlin,         // row (sub) of the finished base-lin angular block -sum_e Cross(f_e) J_pos of a dynamic sample
  kItemDynCol,          // one euler column (sub) of the finished base-ang block of a dynamic sample
  kItemRomFoot,         // range_of_motion_constraint.cc:72-131: 3 g rows + d(R^T r)/d(euler) of one (sample, foot)
  kItemNone = 255,      // padding: keeps every warp of stage A on one item kind
};
constexpr int kItemPad = 16;   // items of one kind are padded to a multiple of this (pack.cc)
// relative cost of an item kind: lower = longer dependent chain (pack.cc orders the kinds of a level by it)
inline int ItemCostRank(int type) {
  switch (type) {
    case kItemDynPre: case kItemDynCol: return 0;
    case kItemForce: case kItemTorque: case kItemForceNode: case kItemRomFoot: case kItemTerrainHard: case kItemTorqueNode: return 1;
    case kItemDynFoot: case kItemRomPre: return 2;
    case kItemDynJlin: case kItemBaseRom: return 3;
    case kItemDynWrench: return 4;
    default: return 5;
  }
}
inline int ItemLevel(int type) { return type >= kItemDynWrench && type != kItemNone ? 1 : 0; }

// scratch layouts written by the level-0 items (slot offsets)
enum { kDpV = 0, kDpA = 3, kDpTrig = 6, kDpR = 12, kDpRI = 21, kDpIw = 30, kDpM = 39, kDpMd = 44, kDpW = 49, kDpWd = 52,
       kDpIww = 55, kDpV11 = 58, kDpV21 = 61, kDynPreSize = 64 };
enum { kRpLin = 0, kRpR = 3, kRpDR = 12, kRomPreSize = 39 };
#ifdef __CUDACC__
#define TB_PHD __host__ __device__
#else
#define TB_PHD
#endif
// G layout of a dynamic sample: finished base-ang block, finished base-lin angular block, f_e[3E], (c - p_e)[3E],
// then scratch: per-foot torque terms [3E], DynPre block
enum { kDynJang = 0, kDynJlin = 36, kDynF = 72 };
// optimised durations: -f_e[3E], -(c - p_e)[3E] follow the fixed-duration block, then one block per foot (ItemC::aux of
// the foot's DynFoot item): W_p | W_f | W_tau | finished ee-schedule block [6][P-1]
TB_PHD inline int DynNegOffset(int E) { return (72 + 9 * E + kDynPreSize) | 1; }
// classes that declare the curvature block (TerrainCurvatureDims): slot offset, within a Force / Torque item, of the
// scales S[dim][rows] (f . d b_i / d(dim)); with optimised durations the ee-motion weights W_p follow them.
// n_w: Hermite-weight slots already there (W_f, or W_tau + W_f), n_sched: ee-schedule columns (0 with fixed durations)
TB_PHD inline int CurvDimCount(int dims) { return (dims & 1) + ((dims >> 1) & 1); }
TB_PHD inline int ForceCurvOffset(int n_w, int n_sched) { return 15 + n_w + 5 * n_sched; }
TB_PHD inline int TorqueCurvOffset(int n_w, int n_sched) { return 15 + n_w + 4 * n_sched; }

constexpr int kMaxEE = 4;
// indices into ItemRec::s
enum { kSBaseLin = 0, kSBaseAng = 1, kSMotion = 2, kSForce = 2 + kMaxEE, kSTorque = 2 + 2 * kMaxEE, kSCount = 2 + 3 * kMaxEE };

struct ItemRec {
  int type;
  int gslot;            // first G slot of this item (segment-relative)
  short ee;             // foot (force / terrain / swing) ; splineacc: unused
  short sub;            // dynamic: 0 = wrench part, 1..3 = euler column (sub-1) of the base-ang block
  short n_ee;
  short pad;
  int aux;              // level-1 items: first scratch slot of their level-0 parent
  int b0;               // dynamic: first of the 16 Hermite basis constants of this sample in B (see structure.cc)
  int row[kMaxEE];      // first constraint row written (absolute in g); RoM: one per foot
  int s[kSCount];       // SampleRec indices; node-based items: node slot / polynomial index in s[0]
};

// one stored Jacobian entry, in storage (row-major) order: G slot (low 16 bits) | B index (bits 16..30);
// bit 31 (kEntryGG): the second factor is a G slot as well (runtime Hermite weights of a PhaseSpline column, classes
// with optimised durations).  The device copy (pack.cc) holds BYTE offsets (slot*8, index*8), the flag in bit 16.
using EntryRec = uint32_t;
constexpr uint32_t kEntryGG = 0x80000000u;

// The device copy of the descriptors (HostProgram::entries_dev) is in storage order too: thread t of stage B takes
// entries t, t + blockDim, ... of its segment, so a warp's descriptor load and its P stores are all coalesced.
constexpr int kEntryThreads = 256;   // threads per problem lane
constexpr uint32_t kNoEntry = 0xFFFFFFFFu;
struct SegmentRec {
  int item0, n_items;   // items of this segment: n_items0 level-0 items first, then the level-1 items
  int n_items0, pad;
  int entry0, n_entries;// entries fed by this segment (contiguous in storage order)
  int g_size;           // G slots used
  int pad2;
};

// BasisCode of an entry whose G factor is a terrain-basis component (constant on a terrain without slope):
//   kCodeForce + 3 i + d   : b_i[d], i = 0..4: n, t1 - mu n, t1 + mu n, t2 - mu n, t2 + mu n   (ItemForce, ItemForceNode)
//   kCodeTorque + 3 i + d  : i = 0..4: t1, t2, n, -n, (-k mu) n                               (ItemTorque)
//   kCodeHeight + d        : -h_x, -h_y                                                       (ItemHeightNode)
enum { kCodeForce = 0, kCodeTorque = 15, kCodeHeight = 30, kCodeCount = 32 };

// slots 0.. of every segment's G buffer
enum { kGOne = 0, kGMass = 1, kGMinusOne = 2, kGReserved = 3 };

// ---- compact class tables: one byte blob per structure class, copied into shared memory once per CTA so that
// stage A never touches global memory (with ~200 KB of shared memory in use per SM the L1 is almost gone and every
// table lookup would be an L2 round trip)
struct ItemC {          // 16 bytes
  uint8_t type, sub, ee, n_ee;
  uint16_t gslot, b0;
  uint32_t row;         // first constraint row
  uint16_t s0;          // first compact sample of the item; node items: node slot; splineacc: polynomial index
  uint16_t aux;         // level-1 items: first scratch slot of the level-0 parent
};
// compact sample = polynomial index | time index << 16 ; compact polynomial = node slot | duration index << 16
// sample order of an item: DynPre [ang], DynFoot [lin, motion, force, torque], DynWrench [lin], RomPre [lin, ang],
// RomFoot [motion], Force [motion, force], Torque [motion, force, torque], BaseRom [lin, ang]
// ---- classes with optimised phase durations (PhaseSpline path, phase_spline.cc / phase_durations.cc): which
// polynomial a sample time falls into, its local time and the Hermite weights are per-problem quantities
struct RtSpline {       // one PhaseSpline: index kind*E + foot, kind 0 = ee-motion, 1 = ee-force, 2 = ee-torque, 3 = ee-ang
                        // (kind 3 only in classes with an EELinearConstraint on the angle splines)
  int poly0, n_polys;   // global polynomial range
  int col0, n_cols;     // its variable set's columns
};
struct RtFoot {         // one PhaseDurations set (16 bytes: one shared-memory load)
  short phase0, n_phases; // offset into the per-problem phase-duration array; P
  int col0;             // first ee-schedule column (P - 1 of them)
  double total;         // t_total_ of THIS foot: its own initial durations summed in order (phase_durations.cc:40-50);
                        // feet may differ from each other in the last bit (2.4 vs 2.3999999999999995)
};
// per global polynomial: phase | polys_in_phase << 8 | prev_polys_in_phase << 16 | foot << 24 ; kPolyFixed for the
// base splines (class-constant duration)
constexpr uint32_t kPolyFixed = 0xFFFFFFFFu;

struct BlobLayout {     // byte offsets into the blob (all 16-byte aligned), element counts
  int bytes;
  int off_B, n_B;                 // double
  int off_times, n_times;         // double[3] per entry: t, pow(t,2), pow(t,3)
  int off_durs, n_durs;           // double[3] per entry: T, pow(T,2), pow(T,3)
  int off_rdurs;                  // double[2] per entry: 1/pow(T,2), 1/pow(T,3) correctly rounded (ExactDiv)
  int off_segments, n_segments;   // SegmentRec
  int off_items, n_items;         // ItemC
  int off_samples, n_samples;     // uint32
  int off_polys, n_polys;         // uint32
  int off_node_src, n_node_values;// uint16, 0xFFFF = pinned to zero
  int off_rt_splines, n_rt_splines;   // RtSpline (3 * n_ee, optimised durations only)
  int off_rt_feet, n_rt_feet;         // RtFoot
  int off_poly_meta;                  // uint32 per polynomial
  int n_pdur;                         // phase durations per problem (sum of n_phases)
  int rt;                             // 1: optimised durations
};

struct DeviceProgram {  // pointers into device memory (one copy per GPU)
  int n_vars, n_rows, nnz, n_ee;
  int n_segments;
  int g_capacity;       // max g_size over segments
  int curvature_dims;   // != 0: the class declares the curvature block (CURV kernel instantiations)
  double total_time;
  BlobLayout blob;
  const unsigned char* blob_data;
  const uint32_t* entries;   // G byte offset | B byte offset << 16, storage order
  const SegmentRec* segments;
};

// ---- shared-memory geometry of the evaluation kernel (eval_kernel.cuh), used by the host-side sizing as well.
// RING variants stage the products of stage B in per-warp rings: kRingDepth slots of kRingBlock = 32 * kRingUB entries
// per warp and problem lane; each lane's row has 2 spare elements (destination parity).
constexpr int kRingDepth = 2;
constexpr int kRingUB = 8;
constexpr int kRingBlock = 32 * kRingUB;
// doubles one stream of `lanes` problem lanes needs next to the class tables (even: keeps 16-byte alignment)
// mode: 0 lock-step, 1 bulk-store ring, 2 split (producer / consumer warp groups: two G buffers)
TB_PHD inline int StreamDoubles(const DeviceProgram& P, int lanes, int mode) {
  const int rt = P.blob.rt ? 8 * P.blob.n_polys + P.blob.n_pdur : 0;   // kRtd (device_math.cuh)
  int n = lanes * (P.blob.n_node_values + 6 * P.blob.n_polys + rt + (mode == 2 ? 2 : 1) * P.g_capacity);
  if (mode == 1) n += (kEntryThreads * lanes / 32) * kRingDepth * lanes * (kRingBlock + 2) + 2;
  return (n + 1) & ~1;
}

struct ProblemConst {   // per-problem constants read by the kernels
  double mass;
  double I[9];          // body inertia tensor, row-major, off-diagonals negated (single_rigid_body_dynamics.cc:36-44)
  int terrain_id;
  int pad;
  TerrainTable table;   // terrain_id == kTerrainTableId: the problem's height profile (terrain.h)
};

// ---- host-side description ------------------------------------------------------------------------------------
struct NodeRef { int node, deriv, dim; };

enum SetKind { kBaseLin = 0, kBaseAng, kEEMotion, kEEAng, kEEForce, kEETorque, kEESchedule };

struct VarSet {         // one ifopt variable set = one column block
  std::string name;
  int kind = 0;
  int ee = -1;
  int col0 = 0, n_vars = 0;
  int n_nodes = 0;
  int node_slot0 = -1;  // first node slot in the device nodes[] array (-1: not staged)
  int poly0 = -1;       // first global polynomial index (-1: no spline staged)
  bool in_nlp = true;   // ee-schedule sets exist always but are columns only when durations are optimised
  std::vector<std::vector<NodeRef>> refs;      // per variable: the node values it writes
  std::vector<int> poly_phase, poly_in_phase, polys_in_phase;
  std::vector<char> poly_const;
  std::vector<double> poly_T;
  int n_polys() const { return n_nodes - 1; }
  bool IsConstantNode(int node) const;
  std::vector<int> NonConstantNodes() const;
  int PolyAtStartOfPhase(int phase) const;
  int PhaseOfNode(int node) const;
  int OptIndex(int node, int deriv, int dim) const;   // nodes_variables.cc:44-54 (-1 if not optimised)
};

struct ConSet { std::string name; int row0, n_rows; };

// what ConstraintSet::GetBounds returns for one row (values depend on per-problem model data)
enum RowBoundKind { kBoundZero = 0, kBoundTerrainSwing, kBoundRom, kBoundForceNormal, kBoundLE0, kBoundGE0,
                    kBoundBaseHeight, kBoundTotalDuration, kBoundNone, kBoundTorqueX, kBoundTorqueY,
                    kBoundTorqueNodeZ,     // +- k mu f_n with the start-of-phase force of the initial guess; dim = phase
                    kBoundEELinear };      // +- tolerance of EELinearConstraint definition `dim`
struct RowBound { int kind, ee, dim; };

// a run of consecutive segments with its own class tables (pack.cc): classes with few samples have one
struct ProgramPart {
  BlobLayout blob{};
  std::vector<unsigned char> blob_data;
  std::vector<SegmentRec> segments;    // item0 relative to this part's items, entry0 absolute
  int g_capacity = 0;                  // max g_size over its segments
};

struct HostProgram {
  int n_vars = 0, n_rows = 0, nnz = 0, n_ee = 0;
  bool optimize_durations = false;
  int curvature_dims = 0;            // TerrainCurvatureDims of the class: the force / torque ee-motion block is declared for these
  double total_time = 0.0;
  std::vector<VarSet> var_sets;      // NLP column order (base-lin, base-ang, ee-motion.., ee-ang.., ee-force.., ee-torque.., [ee-schedule..])
  std::vector<ConSet> con_sets;
  std::vector<RowBound> row_bounds;  // one per constraint row
  std::vector<int> pat_rows, pat_cols;
  std::vector<int> node_src;
  std::vector<PolyRec> polys;
  std::vector<SampleRec> samples;
  std::vector<ItemRec> items;
  std::vector<EntryRec> entries;
  std::vector<double> B;
  std::vector<SegmentRec> segments;
  std::vector<RtSpline> rt_splines;
  std::vector<RtFoot> rt_feet;
  std::vector<uint32_t> poly_meta;
  // packed for the device (pack.cc): one ProgramPart per kernel launch of an evaluation.  blob / blob_data mirror
  // parts[0]; segments is the concatenation of the parts' (padded) segment lists
  std::vector<ProgramPart> parts;
  BlobLayout blob{};
  std::vector<unsigned char> blob_data;
  std::vector<uint32_t> entries_dev;    // byte-offset form of entries (B offsets local to the entry's part)
  // storage positions of the entries whose value depends on x (ascending).  The others are G[kGOne | kGMinusOne | kGMass]
  // * B: constant for a problem (ProblemConstantValues); the host interface can leave them off the bus
  std::vector<int> dyn_index;
  // the same when every problem of a batch is on a terrain without slope (TerrainHasSlope, terrain.h): the terrain basis
  // is then constant as well, and so are the entries tagged with a BasisCode (entry_code, one per stored entry, -1: none)
  std::vector<int> dyn_index_flat;
  std::vector<signed char> entry_code;
  int g_capacity = 0;
  int n_node_values = 0;
  std::string signature;   // canonical text of every structural field (structure_accepts compares it)
  const VarSet* Find(int kind, int ee = -1) const;
};

}  // namespace tb
