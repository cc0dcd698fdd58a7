// sph_simulation over several GPUs of one box (addition: the reference is single-device,
// libclsph/sph_simulation.cpp:273-275).  `devices` > 1 partitions the particles into slabs along
// x, one slab engine per GPU (include/sphb.h, "slab engines"), each driven by its own host thread:
// a step is one sphb_slab_step per rank, issued by all the threads at once — the halo exchange and
// the 8-float all-reduce inside it go over NCCL between the engines' streams.  The frame loop, the
// hooks and save_frame stay those of the single-device path (sph_simulation.cpp); only its three
// primitives — step, read the particles back, upload them — are routed here.
//
// What a hook / save_frame sees: all N particles, stably ordered by grid_index (the reference's
// Morton key under the global grid) with the slabs taken left to right — the reference's
// cell-sorted order up to the order inside a cell, which there depends on the history of the
// sort and is not reproduced across slabs.
#include <algorithm>
#include <cmath>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <limits>
#include <mutex>
#include <thread>
#include <vector>

#include "sph_simulation.h"
#include "sphb.h"

namespace sphb_host {

// n persistent host threads; run(f) executes f(rank) on all of them at once and waits
class worker_team {
 public:
  explicit worker_team(int n) : job_(0), generation_(0), pending_(0), quit_(false), rc_(n, 0) {
    for (int r = 0; r < n; ++r) threads_.push_back(std::thread(&worker_team::loop, this, r));
  }
  ~worker_team() {
    {
      std::lock_guard<std::mutex> lock(m_);
      quit_ = true;
      ++generation_;
    }
    go_.notify_all();
    for (size_t r = 0; r < threads_.size(); ++r) threads_[r].join();
  }
  // Called ON THE WORKER THREAD the moment its job returns non-zero, before anything waits for the
  // other ranks: a rank that fails between two collectives (capacity, a CUDA error, the 20 s halo
  // time-out) leaves its peers inside an all-reduce that can never complete, so run() itself
  // would never return.  The slab driver prints the engine's message there and ends the process.
  std::function<void(int /*rank*/, int /*rc*/, const char* /*what*/)> on_failure;

  // the first non-zero result in rank order (and which rank it came from), else 0
  int run(const std::function<int(int)>& f, int* failed_rank = 0, const char* what = "slab call") {
    std::unique_lock<std::mutex> lock(m_);
    what_ = what;
    job_ = &f;
    pending_ = (int)threads_.size();
    ++generation_;
    go_.notify_all();
    while (pending_ > 0) done_.wait(lock);
    job_ = 0;
    for (size_t r = 0; r < rc_.size(); ++r)
      if (rc_[r] != 0) {
        if (failed_rank) *failed_rank = (int)r;
        return rc_[r];
      }
    return 0;
  }

 private:
  void loop(int rank) {
    unsigned long seen = 0;
    for (;;) {
      const std::function<int(int)>* job = 0;
      {
        std::unique_lock<std::mutex> lock(m_);
        while (generation_ == seen) go_.wait(lock);
        seen = generation_;
        if (quit_) return;
        job = job_;
      }
      const int rc = (*job)(rank);
      if (rc != 0 && on_failure) on_failure(rank, rc, what_);
      {
        std::lock_guard<std::mutex> lock(m_);
        rc_[rank] = rc;
        if (--pending_ == 0) done_.notify_all();
      }
    }
  }
  std::vector<std::thread> threads_;
  std::mutex m_;
  std::condition_variable go_, done_;
  const std::function<int(int)>* job_;
  const char* what_ = "slab call";
  unsigned long generation_;
  int pending_;
  bool quit_;
  std::vector<int> rc_;
};

struct slab_team {
  explicit slab_team(int n)
      : world(n), workers(n), engines(n, (sphb_engine*)0), stage(0), stage_ids(0), hist(n), owned(n, 0),
        capacity(0), exchange_capacity(0) {}
  int world;
  worker_team workers;
  std::vector<sphb_engine*> engines;
  std::vector<float> cuts;           // world + 1 slab faces, -inf ... +inf
  std::vector<uint32_t> ids;         // id of the particle at each index of the caller's host array
  particle* stage;                   // particles in rank order (uploads and downloads), page-locked
  uint32_t* stage_ids;               // (so that every rank's copy runs at full PCIe rate, all at once)
  std::vector<std::vector<uint32_t> > hist;  // per rank: particles per grid_index, then their offsets
  std::vector<uint32_t> owned;       // particles each rank owns
  std::vector<particle> recut;       // a re-cut's own gather buffer (the caller's array may be on loan to save_frame)
  uint32_t capacity, exchange_capacity;
};

// Slab faces for `world` ranks holding (nearly) equal counts: a face lies midway between two
// neighbouring DISTINCT x values, so that a lattice plane is never split; the outer faces are
// -inf / +inf.  (libclsph_b200/slab.py::equal_count_cuts is the same rule.)
std::vector<float> equal_count_cuts(const particle* p, size_t n, int world) {
  std::vector<float> xs(n);
  for (size_t i = 0; i < n; ++i) xs[i] = p[i].position.s[0];  // (NaN positions are not expected here)
  std::vector<float> cuts;
  cuts.push_back(-std::numeric_limits<float>::infinity());
  // only world - 1 order statistics are needed, not the sorted array: each selection works on
  // the part of the array right of the previous one
  size_t done = 0;  // xs[0 .. done) already hold the `done` smallest values
  for (int r = 1; r < world; ++r) {
    const size_t k = n * (size_t)r / (size_t)world;
    float lo = 0.f, hi = 0.f;
    if (n) {
      const size_t at = k > 0 ? k - 1 : 0;  // lo = the (at + 1)-th smallest value
      if (at >= done) {
        std::nth_element(xs.begin() + done, xs.begin() + at, xs.end());
        done = at + 1;
      }
      lo = xs[at];
      // hi = the smallest value strictly greater than lo (everything left of `at` is <= lo)
      bool found = false;
      for (size_t i = at + 1; i < n; ++i)
        if (xs[i] > lo && (!found || xs[i] < hi)) {
          hi = xs[i];
          found = true;
        }
      if (!found) hi = lo;
    }
    cuts.push_back((float)(((double)lo + (double)hi) / 2));
  }
  cuts.push_back(std::numeric_limits<float>::infinity());
  return cuts;
}

// which slab [cuts[r], cuts[r+1]) holds x
inline int slab_of(const std::vector<float>& cuts, float x) {
  const int world = (int)cuts.size() - 1;
  int r = (int)(std::upper_bound(cuts.begin() + 1, cuts.end() - 1, x) - (cuts.begin() + 1));
  return r < 0 ? 0 : (r >= world ? world - 1 : r);
}

// Stable counting sort by grid_index — slabs left to right (stage[start[r] .. start[r+1]) is slab
// r), each in its own order — into `particles` / `ids_out`, by all the host threads at once.
void gather_sort(worker_team& workers, std::vector<std::vector<uint32_t> >& hist, const particle* stage,
                 const uint32_t* stage_ids, const std::vector<uint32_t>& start, uint32_t cells,
                 particle* particles, uint32_t* ids_out) {
  const int world = (int)start.size() - 1;
  const uint32_t n = start[world];
  if ((int)hist.size() < world) hist.resize(world);
  // keys past the grid's Morton bound (runaway particles) share one last bucket
  if ((size_t)cells + 1 > ((size_t)1 << 26)) {  // a table per rank would be too much: one thread, one table
    std::vector<uint32_t> first((size_t)cells + 2, 0);
    for (uint32_t i = 0; i < n; ++i) {
      const uint32_t k = stage[i].grid_index < cells ? stage[i].grid_index : cells;
      ++first[(size_t)k + 1];
    }
    for (size_t c = 0; c <= cells; ++c) first[c + 1] += first[c];
    for (uint32_t i = 0; i < n; ++i) {
      const uint32_t k = stage[i].grid_index < cells ? stage[i].grid_index : cells;
      const uint32_t d = first[k]++;
      particles[d] = stage[i];
      ids_out[d] = stage_ids[i];
    }
    return;
  }
  // The same by all the host threads at once: every rank counts its slab's keys, the keys are
  // split into one range per thread for the prefix over (key, rank), every rank scatters its slab.
  const size_t buckets = (size_t)cells + 1;
  const int W = world;
  std::vector<uint64_t> range_total(W, 0), range_base(W + 1, 0);
  workers.run([&](int r) -> int {
    std::vector<uint32_t>& h = hist[r];
    h.assign(buckets, 0u);
    for (uint32_t i = start[r]; i < start[r + 1]; ++i) {
      const uint32_t k = stage[i].grid_index < cells ? stage[i].grid_index : cells;
      ++h[k];
    }
    return 0;
  });
  workers.run([&](int w) -> int {
    const size_t lo = buckets * (size_t)w / W, hi = buckets * (size_t)(w + 1) / W;
    uint64_t sum = 0;
    for (int r = 0; r < W; ++r) {
      const std::vector<uint32_t>& h = hist[r];
      for (size_t k = lo; k < hi; ++k) sum += h[k];
    }
    range_total[w] = sum;
    return 0;
  });
  for (int w = 0; w < W; ++w) range_base[w + 1] = range_base[w] + range_total[w];
  workers.run([&](int w) -> int {
    const size_t lo = buckets * (size_t)w / W, hi = buckets * (size_t)(w + 1) / W;
    uint32_t run = (uint32_t)range_base[w];
    for (size_t k = lo; k < hi; ++k)
      for (int r = 0; r < W; ++r) {  // within a key: slabs left to right
        const uint32_t cnt = hist[r][k];
        hist[r][k] = run;
        run += cnt;
      }
    return 0;
  });
  workers.run([&](int r) -> int {
    std::vector<uint32_t>& h = hist[r];
    for (uint32_t i = start[r]; i < start[r + 1]; ++i) {
      const uint32_t k = stage[i].grid_index < cells ? stage[i].grid_index : cells;
      const uint32_t d = h[k]++;
      particles[d] = stage[i];
      ids_out[d] = stage_ids[i];
    }
    return 0;
  });
}

// The opposite direction: the caller's array dealt out to the slabs [cuts[r], cuts[r+1]) —
// stage[start[r] .. start[r+1]) receives slab r's particles in the order they have in the
// caller's array — by all the host threads at once, each owning one contiguous piece of the input.
void deal_out(worker_team& workers, int threads, const std::vector<float>& cuts, const particle* particles,
              const uint32_t* ids, uint32_t n, particle* stage, uint32_t* stage_ids, std::vector<uint32_t>& start) {
  const int world = (int)cuts.size() - 1;
  const int W = threads < 1 ? 1 : threads;
  std::vector<unsigned char> where(n);  // (a communicator has at most 255 ranks here: 8 in practice)
  std::vector<std::vector<uint32_t> > count(W, std::vector<uint32_t>(world, 0u));
  workers.run([&](int w) -> int {
    if (w >= W) return 0;
    const size_t lo = (size_t)n * w / W, hi = (size_t)n * (w + 1) / W;
    std::vector<uint32_t>& c = count[w];
    for (size_t i = lo; i < hi; ++i) {
      const int r = slab_of(cuts, particles[i].position.s[0]);
      where[i] = (unsigned char)r;
      ++c[r];
    }
    return 0;
  });
  start.assign(world + 1, 0u);
  uint32_t run = 0;
  for (int r = 0; r < world; ++r) {  // slab by slab, pieces of the input in order
    start[r] = run;
    for (int w = 0; w < W; ++w) {
      const uint32_t c = count[w][r];
      count[w][r] = run;
      run += c;
    }
  }
  start[world] = run;
  workers.run([&](int w) -> int {
    if (w >= W) return 0;
    const size_t lo = (size_t)n * w / W, hi = (size_t)n * (w + 1) / W;
    std::vector<uint32_t>& cursor = count[w];
    for (size_t i = lo; i < hi; ++i) {
      const uint32_t d = cursor[where[i]]++;
      stage[d] = particles[i];
      stage_ids[d] = ids[i];
    }
    return 0;
  });
}

// (tests: the two directions of a frame without any engine)
int deal_out_selftest(const particle* particles, const uint32_t* ids, uint32_t n, const float* cuts, int world,
                      int threads, particle* stage, uint32_t* stage_ids, uint32_t* start_out) {
  if (world < 1 || world > 255 || threads < 1) return -1;
  worker_team workers(threads);
  const std::vector<float> c(cuts, cuts + world + 1);
  std::vector<uint32_t> start;
  deal_out(workers, threads, c, particles, ids, n, stage, stage_ids, start);
  for (int r = 0; r <= world; ++r) start_out[r] = start[r];
  return 0;
}

int gather_sort_selftest(const particle* stage, const uint32_t* stage_ids, const uint32_t* start, int world,
                         uint32_t cells, particle* particles, uint32_t* ids_out) {
  worker_team workers(world);
  std::vector<std::vector<uint32_t> > hist(world);
  const std::vector<uint32_t> st(start, start + world + 1);
  gather_sort(workers, hist, stage, stage_ids, st, cells, particles, ids_out);
  return 0;
}

}  // namespace sphb_host

using sphb_host::slab_team;

void sph_simulation::slabs_fail(int rc, int rank, const char* what) {
  std::cerr << "A CUDA engine error occured (" << what << ", GPU rank " << rank << ")-> "
            << ((team_ && team_->engines[rank]) ? sphb_last_error(team_->engines[rank]) : sphb_error_string(rc))
            << std::endl;
  std::cerr.flush();
  std::cout.flush();
  // The reference's convention is a message and exit(-1) (util/cl_boilerplate.h:32-34).  _Exit and
  // not exit: the other ranks' threads may sit inside a collective that will never complete, and a
  // normal exit would wait for them while tearing down the CUDA runtime.
  std::_Exit(255);
}

// Engines, scene, communicator.  Every rank gets room for 1.6 x its fair share plus two ghost
// layers; a halo message holds up to N / 4 records (at least 65 536).
void sph_simulation::slabs_begin() {
  const int world = devices;
  const uint32_t n = parameters.particles_count;
  if (world > 255) {  // (deal_out keeps a particle's slab in one byte)
    std::cerr << "sph_simulation::devices: at most 255 GPUs" << std::endl;
    std::exit(-1);
  }
  team_ = new slab_team(world);
  slab_team& T = *team_;
  uint64_t exch = n / 4u;
  if (exch < 65536u) exch = 65536u;
  if (exch > n) exch = n ? n : 1u;
  uint64_t cap = (uint64_t)(1.6 * (double)n / world) + 2u * exch + 1024u;
  if (cap > n) cap = n ? n : 1u;
  T.capacity = (uint32_t)cap;
  T.exchange_capacity = (uint32_t)exch;
  T.ids.resize(n);
  for (uint32_t i = 0; i < n; ++i) T.ids[i] = i;
  check(sphb_host_alloc(reinterpret_cast<void**>(&T.stage), sizeof(particle) * (size_t)(n ? n : 1)), __FILE__, __LINE__);
  check(sphb_host_alloc(reinterpret_cast<void**>(&T.stage_ids), sizeof(uint32_t) * (size_t)(n ? n : 1)), __FILE__,
        __LINE__);

  unsigned char id[128];
  int rc = sphb_nccl_unique_id(id);
  if (rc != 0) {
    std::cerr << "A CUDA engine error occured (NCCL)-> " << sphb_last_error(0) << std::endl;
    std::exit(-1);
  }
  const int base = device < 0 ? 0 : device;
  const sphb_params* P = reinterpret_cast<const sphb_params*>(&parameters);
  const sphb_terms* K = reinterpret_cast<const sphb_terms*>(&precomputed_terms);
  int bad = 0;
  rc = T.workers.run([&](int r) -> int {
    sphb_engine* e = 0;
    if (int c = sphb_slab_create(P, K, base + r, T.capacity, T.exchange_capacity, &e)) return c;
    T.engines[r] = e;
    return sphb_set_scene(e, reinterpret_cast<const sphb_bb*>(current_scene.bbs.data()),
                          current_scene.face_count, current_scene.transforms.data(),
                          current_scene.rvertices.data(), current_scene.totalGridpoints);
  }, &bad);
  if (rc != 0) slabs_fail(rc, bad, "slab set-up");
  // from here on the ranks meet in collectives: a rank that fails reports and ends the process at once
  T.workers.on_failure = [this](int r, int c, const char* what) { slabs_fail(c, r, what); };
  // (collective: entered only once every rank has an engine, so a missing GPU is an error
  // message above and not a rendezvous that never completes)
  rc = T.workers.run([&](int r) -> int { return sphb_slab_comm_init(T.engines[r], id, r, world); }, &bad,
                     "NCCL communicator");
  if (rc != 0) slabs_fail(rc, bad, "NCCL communicator");
}

void sph_simulation::slabs_end() {
  if (!team_) return;
  slab_team& T = *team_;
  T.workers.run([&](int r) -> int {
    if (T.engines[r]) sphb_destroy(T.engines[r]);
    T.engines[r] = 0;
    return 0;
  });
  sphb_host_free(T.stage);
  sphb_host_free(T.stage_ids);
  delete team_;
  team_ = 0;
}

// The caller's host array -> the slab engines.  `recut`: new equal-count faces from the current
// positions (else the particles are dealt out by the faces in force).
void sph_simulation::slabs_upload(const particle* particles, bool recut) {
  slab_team& T = *team_;
  const uint32_t n = parameters.particles_count;
  if (recut || T.cuts.empty()) T.cuts = sphb_host::equal_count_cuts(particles, n, T.world);
  std::vector<uint32_t> start;
  sphb_host::deal_out(T.workers, T.world, T.cuts, particles, T.ids.data(), n, T.stage, T.stage_ids, start);
  for (int r = 0; r < T.world; ++r) T.owned[r] = start[r + 1] - start[r];
  int bad = 0;
  const int rc = T.workers.run([&](int r) -> int {
    sphb_engine* e = T.engines[r];
    if (int c = sphb_slab_configure(e, T.cuts[r], T.cuts[r + 1], r > 0, r + 1 < T.world)) return c;
    if (int c = sphb_slab_upload(e, reinterpret_cast<const sphb_particle*>(T.stage + start[r]),
                                 T.stage_ids + start[r], T.owned[r]))
      return c;
    return sphb_slab_start(e);  // collective: the global bbox
  }, &bad);
  if (rc != 0) slabs_fail(rc, bad, "slab upload");
}

// The slab engines -> the caller's host array, ordered by grid_index (see the file header).
void sph_simulation::slabs_read(particle* particles) {
  slab_team& T = *team_;
  const uint32_t n = parameters.particles_count;
  int bad = 0;
  int rc = T.workers.run([&](int r) -> int {
    uint32_t cnt = 0;
    const int c = sphb_slab_download(T.engines[r], 0, 0, 0, &cnt);
    T.owned[r] = cnt;
    return c;
  }, &bad);
  if (rc != 0) slabs_fail(rc, bad, "slab download");
  std::vector<uint32_t> start(T.world + 1, 0);
  for (int r = 0; r < T.world; ++r) start[r + 1] = start[r] + T.owned[r];
  if (start[T.world] != n) {
    std::cerr << "slab engines hold " << start[T.world] << " particles, expected " << n << std::endl;
    std::exit(-1);
  }
  rc = T.workers.run([&](int r) -> int {
    uint32_t cnt = 0;
    if (!T.owned[r]) return 0;
    return sphb_slab_download(T.engines[r], reinterpret_cast<sphb_particle*>(T.stage + start[r]),
                              T.stage_ids + start[r], T.owned[r], &cnt);
  }, &bad);
  if (rc != 0) slabs_fail(rc, bad, "slab download");
  sphb_host::gather_sort(T.workers, T.hist, T.stage, T.stage_ids, start, parameters.grid_cell_count, particles,
                         T.ids.data());
}

// One simulate_single_frame over all slabs; afterwards, when the fullest slab has outgrown
// `rebalance` x the mean, the particles are gathered and dealt out again by fresh equal-count faces.
float sph_simulation::slabs_step(float dt) {
  slab_team& T = *team_;
  std::vector<float> out(T.world, dt);
  int bad = 0;
  const int rc = T.workers.run([&](int r) -> int {
    int attempts = 0;
    uint32_t ghosts = 0;
    return sphb_slab_step(T.engines[r], dt, &out[r], &attempts, &T.owned[r], &ghosts);
  }, &bad, "slab step");
  if (rc != 0) slabs_fail(rc, bad, "slab step");
  for (int r = 1; r < T.world; ++r)
    if (std::memcmp(&out[r], &out[0], sizeof(float)) != 0) {
      std::cerr << "slab ranks disagree on the time step (" << out[0] << " vs " << out[r] << ")" << std::endl;
      std::exit(-1);
    }
  sphb_params now;
  check(sphb_get_parameters(T.engines[0], &now), __FILE__, __LINE__);
  parameters.grid_size_x = now.grid_size_x;
  parameters.grid_size_y = now.grid_size_y;
  parameters.grid_size_z = now.grid_size_z;
  parameters.grid_cell_count = now.grid_cell_count;
  for (int a = 0; a < 4; ++a) {
    parameters.min_point.s[a] = now.min_point.s[a];
    parameters.max_point.s[a] = now.max_point.s[a];
  }
  ++steps_done;
  if (rebalance > 0.f && (T.world > 1 || force_slabs)) {
    uint32_t most = 0;
    for (int r = 0; r < T.world; ++r) most = std::max(most, T.owned[r]);
    if ((double)most > (double)rebalance * parameters.particles_count / T.world) {
      // (sphb_slab_step's counts are those before the step's own migration: one step late, which
      // the capacity head-room covers)
      T.recut.resize(parameters.particles_count);
      slabs_read(T.recut.data());
      slabs_upload(T.recut.data(), true);
      ++slab_recuts;
    }
  }
  return out[0];
}
