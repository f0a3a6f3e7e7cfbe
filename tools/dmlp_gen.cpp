// Native writer for the hashed synthetic stream (text grammar of the reference,
// reference: generate_input.py:11-21).  Usage:
//   dmlp-gen N Q A lo hi minK maxK labels seed output
#include <cstdio>
#include <cstdlib>

#include "dmlp/io.h"

int main(int argc, char **argv) {
  if (argc != 11) {
    std::fprintf(stderr, "usage: %s N Q A lo hi minK maxK labels seed output\n", argv[0]);
    return 2;
  }
  dmlp::SynthSpec s;
  s.n = std::atoll(argv[1]); s.q = std::atoll(argv[2]); s.a = std::atoi(argv[3]);
  s.lo = std::atof(argv[4]); s.hi = std::atof(argv[5]);
  s.min_k = std::atoi(argv[6]); s.max_k = std::atoi(argv[7]);
  s.num_labels = std::atoi(argv[8]); s.seed = std::strtoull(argv[9], nullptr, 10);
  dmlp::Problem p;
  dmlp::synth_problem(s, &p);
  std::FILE *f = std::fopen(argv[10], "w");
  if (!f) { std::perror(argv[10]); return 1; }
  dmlp::write_problem_text(p, f);
  std::fclose(f);
  return 0;
}
