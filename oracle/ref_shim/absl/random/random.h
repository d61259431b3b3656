// ORACLE BUILD SHIM (test infrastructure): absl::BitGen / Gaussian / Bernoulli with an optional REPLAY source.
//
// The reference draws its knot noise inside SamplingPlanner::AddNoiseToPolicy (planners/sampling/planner.cc:313-339): a
// local BitGen per candidate, one Bernoulli(0.2) when a second std is configured, then one Gaussian(0, scale*std) per
// knot value in node-major order.  For parity runs the harness injects the SAME standard-normal tensor z[N][P*nu] (and
// second-std flags) that the CUDA engine and the oracle consume: ref_shim_random::begin_replay() arms it, every BitGen
// that DRAWS afterwards takes the next candidate index at its first draw (the planner runs on a one-thread pool, so draw
// order is candidate order 1, 2, ...; candidate 0 is never noised, and the BitGen Trajectory::NoisyRollout constructs
// never draws while xfrc_std == 0), and Gaussian(mean, sd) returns mean + sd * z.  Without a replay
// source BitGen is a std::mt19937_64 seeded from std::random_device, i.e. the reference's run-to-run behaviour.
#ifndef ORACLE_SHIM_ABSL_RANDOM_H_
#define ORACLE_SHIM_ABSL_RANDOM_H_
#include <random>
namespace ref_shim_random {
struct Replay { const double* z = nullptr; const unsigned char* std2 = nullptr; long long per_candidate = 0, n = 0, next = 1; bool on = false; };
Replay& replay();
inline void begin_replay(const double* z, const unsigned char* std2, long long n, long long per_candidate) {
  Replay& r = replay(); r.z = z; r.std2 = std2; r.n = n; r.per_candidate = per_candidate; r.next = 1; r.on = z != nullptr;
}
inline void end_replay() { replay().on = false; }
}  // namespace ref_shim_random
namespace absl {
class BitGen {
 public:
  BitGen();
  using result_type = std::mt19937_64::result_type;
  static constexpr result_type min() { return std::mt19937_64::min(); }
  static constexpr result_type max() { return std::mt19937_64::max(); }
  result_type operator()() { return eng_(); }
  long long candidate = -1, cursor = 0;     // replay position
  std::mt19937_64 eng_;
};
template <class T, class G> T Gaussian(G& g, T mean = 0, T stddev = 1) {
  ref_shim_random::Replay& r = ref_shim_random::replay();
  if (r.on && g.candidate < 0) g.candidate = r.next++;
  if (r.on && g.candidate >= 0 && g.candidate < r.n && g.cursor < r.per_candidate) return mean + stddev * (T)r.z[g.candidate * r.per_candidate + g.cursor++];
  std::normal_distribution<double> d(0.0, 1.0); return mean + stddev * (T)d(g.eng_);
}
template <class G> bool Bernoulli(G& g, double p) {
  ref_shim_random::Replay& r = ref_shim_random::replay();
  if (r.on && g.candidate < 0) g.candidate = r.next++;
  if (r.on && g.candidate >= 0 && g.candidate < r.n) return r.std2 ? r.std2[g.candidate] != 0 : false;
  std::bernoulli_distribution d(p); return d(g.eng_);
}
template <class T, class G> T Uniform(G& g, T lo, T hi) { std::uniform_real_distribution<double> d(lo, hi); return (T)d(g.eng_); }
}  // namespace absl
#endif
