// ctba_shard.h — time-segment sharding of ONE problem over the GPUs of a node.
//
// Every rank is handed the complete problem through the same C-ABI calls as a single-GPU run (drop-in: the caller
// only adds ctba_comm_init_*).  On the first solve the library
//   1. analyses the global structure on the host (node order, band width, which unknowns are touched),
//   2. cuts the block columns of the reduced system into P = (partitions per rank) x world time partitions
//      (the same interior/separator layout the single-GPU band factorisation uses, ctba_band.cu),
//   3. keeps only the measurements ("units") whose FIRST touched block column lies in this rank's range:
//      a landmark with all its observations, an IMU / GP / bias-cost sample.  A unit spans at most the band
//      width, so the partial normal equations of neighbouring ranks overlap only in a separator's diagonal
//      blocks — and those only ever enter the interface system, which is summed with ONE all-reduce.
// Per LM iteration the exchange steps are: all-reduce of the per-frame squared reprojection norms (a frame in
// the overlap has observations on two ranks and CauchyLoss acts on the frame block, bspline_so3r3.h:1032),
// all-gather of the 8 cost terms, all-reduce of [gradient | diag(J^T J)], all-reduce of the interface system,
// all-gather of the 8 step scalars; the state is re-assembled once at the end of the solve.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>
#include <vector>

namespace ctba {

struct Comm {
  int rank = 0, world = 1;
  virtual ~Comm() {}
  virtual int allreduce_sum(double* buf, size_t n, cudaStream_t st) = 0;                       // in place
  virtual int allgather(const double* send, double* recv, size_t n, cudaStream_t st) = 0;      // recv: world * n
  virtual const char* kind() const = 0;
  virtual const char* error() const { return ""; }
};

// time partitions of the block columns: interior I_q = [c0[q], c0[q] + nI[q]), separator S_q = the nbw columns after it (q < P-1)
struct PartLayout {
  int P = 1;
  std::vector<int> c0, nI;
  int sep_start(int q) const { return q < 0 ? 0 : c0[q] + nI[q]; }   // first block column of S_q (q = P-1: end of the band)
};
// requested: 0 = automatic; world > 1: P is a multiple of world.  Returns 0, or -1 when the band is too short for `world` partitions.
int band_partition_layout(int nbc, int nbw, int requested, int world, int nested, PartLayout* out);

struct ShardInfo {
  int world = 1, rank = 0;
  PartLayout layout;
  int nbw = 0, nbc = 0;
  int q_lo = 0, q_hi = 1;          // partitions of this rank
  int bc_lo = 0, bc_hi = 0;        // block columns this rank's partial system can touch (left separator .. right separator)
  int own_lo = 0, own_hi = 0;      // scalar columns this rank owns (LM damping, step scalars, final state)
  int F_global = 0, f_first = 0, f_cost_begin = 0;
  std::vector<int32_t> lm_owner;   // per landmark: owning rank (unobserved: 0)
  std::vector<int32_t> imu_owner, gp_owner, bias_owner;   // per IMU sample / GP sample / bias random-walk block of the WHOLE problem
  int rank_of_block(int bc) const {   // rank whose partitions' [left separator start, right separator start) holds block column bc
    const int per = layout.P / world;
    for (int q = 0; q < layout.P - 1; ++q) if (bc < layout.sep_start(q)) return q / per;
    return world - 1;
  }
};

// what has to be identical on every rank although the units are filtered
struct GlobalStructure {
  bool has_imu = false, any_frames = false, any_imu = false, any_gp = false, gravity_norm = false;
  int hbn = 0;
  std::vector<uint8_t> tk, tb;
};

struct HostStructure;
}  // namespace ctba
struct ctba_problem;
namespace ctba {
int shard_plan(const ctba_problem* p, int world, int rank, ShardInfo* sh, GlobalStructure* gs, HostStructure* hs_out);
Comm* make_local_comms(int world, std::vector<Comm*>* out);   // in-process group (one device, one host thread per rank)
Comm* make_nccl_comm(int world, int rank, const char* unique_id128, int device, char* err, size_t errlen);
int nccl_unique_id(char* out128, char* err, size_t errlen);

}  // namespace ctba
