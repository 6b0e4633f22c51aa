// drivers/harness.h — what the command-line drivers of the B200 build share: a phase stopwatch that
// prints the reference drivers' report lines (spmv_test.cu:36-40, spmm_test.cu:44-48), the
// all-replicas check (every GPU must hold the complete result, matrix.h:618-622 / :714-726) and the
// argument helpers.  Header-only, plain C++.
#ifndef SBLAS_DRIVERS_HARNESS_H
#define SBLAS_DRIVERS_HARNESS_H

#include <cstdlib>
#include <iostream>
#include <string>

namespace drv {

// Three phases, timed with the reference's cpu_timer (utility.h:130-150) so that the numbers mean
// the same thing as in the reference's reports.
template <typename Timer>
class Phases {
 public:
  enum Phase { LOAD = 0, GPU = 1, CPU = 2 };
  template <typename F>
  void run(Phase p, F&& body) {
    t_[p].start_timer();
    body();
    t_[p].stop_timer();
  }
  // "Validation = ..." first: tests/test_dropin.py greps for it
  void report(bool ok, unsigned n_gpu, bool gpu_line_first = false) {
    std::cout << "Validation = " << (ok ? "True" : "False") << std::endl;
    std::cout << "Load Time: " << t_[LOAD].measure() << "ms." << std::endl;
    if (gpu_line_first) std::cout << n_gpu << "-GPUs Run Time: " << t_[GPU].measure() << " ms." << std::endl;
    std::cout << "CPU Run Time: " << t_[CPU].measure() << " ms." << std::endl;
    if (!gpu_line_first) std::cout << n_gpu << "-GPUs Run Time: " << t_[GPU].measure() << " ms." << std::endl;
  }

 private:
  Timer t_[3];
};

struct Args {
  int argc;
  char** argv;
  const char* str(int i) const { return argv[i]; }
  double real(int i) const { return std::atof(argv[i]); }
  unsigned count(int i) const { return (unsigned)std::atoi(argv[i]); }
  bool flag(int i) const { return i < argc && std::atoi(argv[i]) == 1; }
  int usage(const std::string& line) const {
    std::cerr << line << std::endl;
    return 1;
  }
};

}  // namespace drv

#endif  // SBLAS_DRIVERS_HARNESS_H
