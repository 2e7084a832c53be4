// h2ok-bench — the `h2ok bench` sub-command as a stand-alone binary.
//
// The reference is a Rust CLI (crate `h2ok`); no Rust toolchain exists on this machine, so the sub-command is
// delivered twice: as unverified Rust source under reference_patch/ (how it would be wired into src/cli/mod.rs and
// src/main.rs) and as this C++ program, which links libh2ok_hist.so, takes the same flags and behaves like the
// reference's commands do:
//   * flags are declared in one table with a validator each, like build_app() (src/cli/mod.rs:162-232) and its
//     validators (src/cli/mod.rs:245-279): a bad value prints "error: Invalid value for '--x <x>': ..." and exits 1;
//   * -h / --help prints clap-style help that starts with "h2ok-bench" (the integration tests match sub-command help
//     with r"h2ok-deploy.*" / r"h2ok-undeploy.*", tests/integration_tests.rs:32,45-46);
//   * on a terminal it prints a human-readable table, otherwise exactly one machine-readable line
//     (src/main.rs:62-68 prints the descriptor name only when stdout is a pipe);
//   * the process is single-threaded and blocking (src/main.rs:18-37): all GPUs of --gpus are driven through
//     h2ok_hist_group_* from this one thread.
// No torch, no Python: only the C ABI of include/h2ok_hist.h.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <chrono>
#include <string>
#include <vector>

#include "../../include/h2ok_hist.h"

namespace {

const char* kAppVersion = "0.1.0";  // APP_VERSION, src/cli/mod.rs:14

struct Flag {
  const char* name;   // --name
  char short_name;    // -s, 0: none
  const char* value;  // default
  const char* help;
  bool (*ok)(const char*, std::string*);
};

bool int_gt_zero(const char* v, std::string* why) {  // validate_int_greater_than_zero, src/cli/mod.rs:245-252
  char* end = nullptr;
  long long n = strtoll(v, &end, 10);
  if (!*v || *end) { *why = "invalid digit found in string"; return false; }
  if (n < 1) { *why = "Error: The number provided must be greater than zero."; return false; }
  return true;
}
bool shape_ok(const char* v, std::string* why) {
  if (!strcmp(v, "c2") || !strcmp(v, "c3") || !strcmp(v, "c4")) return true;
  *why = "Error: The shape must be one of c2, c3, c4.";
  return false;
}
bool uint_ok(const char* v, std::string* why) {
  char* end = nullptr;
  long long n = strtoll(v, &end, 10);
  if (!*v || *end || n < 0) { *why = "Error: The number provided must not be negative."; return false; }
  return true;
}

Flag kFlags[] = {
    {"shape", 0, "c3", "Synthetic data shape: c2 (28 HIGGS-shaped columns), c3 (wide, 64 columns) or c4 (sparse).", shape_ok},
    {"rows", 'r', "0", "Number of rows over all GPUs. 0: the shape's own (1M / 100M / 10M).", uint_ok},
    {"cols", 0, "0", "Number of columns. 0: the shape's own (28 / 64 / 2000).", uint_ok},
    {"bins", 'b', "0", "Histogram bins per column (nbins). 0: the shape's own (256 / 256 / 64).", uint_ok},
    {"depth", 'd', "0", "Tree depth (levels per tree). 0: the shape's own (6 / 8 / 6).", uint_ok},
    {"gpus", 'g', "1", "Number of GPUs of this node to shard the rows over.", int_gt_zero},
    {"trees", 't', "3", "Trees to time (one more is built first as a warm-up).", int_gt_zero},
    {"seed", 0, "20261106", "Seed of the counter-based generator.", uint_ok},
};
const int kNumFlags = sizeof(kFlags) / sizeof(kFlags[0]);

void print_help(FILE* f) {
  fprintf(f, "h2ok-bench %s\n", kAppVersion);
  fprintf(f, "Benchmarks the GBM histogram path (histogram build, split search, row routing per tree level) on the GPUs of "
             "this node. Prints a table on a terminal, one JSON line otherwise.\n\n");
  fprintf(f, "USAGE:\n    h2ok bench [OPTIONS]\n\nFLAGS:\n    -h, --help       Prints help information\n"
             "    -V, --version    Prints version information\n\nOPTIONS:\n");
  for (int i = 0; i < kNumFlags; i++) {
    char left[64];
    if (kFlags[i].short_name) snprintf(left, sizeof(left), "-%c, --%s <%s>", kFlags[i].short_name, kFlags[i].name, kFlags[i].name);
    else snprintf(left, sizeof(left), "    --%s <%s>", kFlags[i].name, kFlags[i].name);
    fprintf(f, "    %-28s %s [default: %s]\n", left, kFlags[i].help, kFlags[i].value);
  }
}

const char* flag(const char* name) {
  for (int i = 0; i < kNumFlags; i++)
    if (!strcmp(kFlags[i].name, name)) return kFlags[i].value;
  return "";
}

#define OK(call)                                                        \
  do {                                                                  \
    if ((call) != H2OK_OK) {                                            \
      fprintf(stderr, "h2ok-bench: %s: %s\n", #call, h2ok_last_error()); \
      return 1;                                                         \
    }                                                                   \
  } while (0)

}  // namespace

int main(int argc, char** argv) {
  std::vector<std::string> keep(argc);  // values outlive the loop
  for (int a = 1; a < argc; a++) {
    const char* s = argv[a];
    if (!strcmp(s, "-h") || !strcmp(s, "--help")) { print_help(stdout); return 0; }
    if (!strcmp(s, "-V") || !strcmp(s, "--version")) { printf("h2ok-bench %s\n", kAppVersion); return 0; }
    int hit = -1;
    const char* val = nullptr;
    for (int i = 0; i < kNumFlags && hit < 0; i++) {
      const size_t n = strlen(kFlags[i].name);
      if (s[0] == '-' && s[1] == '-' && !strncmp(s + 2, kFlags[i].name, n) && (s[2 + n] == 0 || s[2 + n] == '=')) {
        hit = i;
        val = s[2 + n] == '=' ? s + 3 + n : nullptr;
      } else if (s[0] == '-' && s[1] == kFlags[i].short_name && kFlags[i].short_name && s[2] == 0) {
        hit = i;
      }
    }
    if (hit < 0) {
      fprintf(stderr, "error: Found argument '%s' which wasn't expected, or isn't valid in this context\n\nUSAGE:\n    h2ok bench [OPTIONS]\n\n"
                      "For more information try --help\n", s);
      return 1;
    }
    if (!val) {
      if (a + 1 >= argc) {
        fprintf(stderr, "error: The argument '--%s <%s>' requires a value but none was supplied\n", kFlags[hit].name, kFlags[hit].name);
        return 1;
      }
      val = argv[++a];
    }
    std::string why;
    if (!kFlags[hit].ok(val, &why)) {
      fprintf(stderr, "error: Invalid value for '--%s <%s>': %s\n", kFlags[hit].name, kFlags[hit].name, why.c_str());
      return 1;
    }
    keep[a] = val;
    kFlags[hit].value = keep[a].c_str();
  }

  const std::string shape = flag("shape");
  const int shape_id = shape == "c2" ? H2OK_SHAPE_HIGGS28 : (shape == "c3" ? H2OK_SHAPE_WIDE64 : H2OK_SHAPE_SPARSE);
  long long rows = atoll(flag("rows")), cols = atoll(flag("cols")), bins = atoll(flag("bins")), depth = atoll(flag("depth"));
  if (!rows) rows = shape == "c2" ? 1000000 : (shape == "c3" ? 100000000 : 10000000);
  if (!cols) cols = shape == "c2" ? 28 : (shape == "c3" ? 64 : 2000);
  if (!bins) bins = shape == "c4" ? 64 : 256;
  if (!depth) depth = shape == "c3" ? 8 : 6;
  const int gpus = atoi(flag("gpus")), trees = atoi(flag("trees"));
  const uint64_t seed = strtoull(flag("seed"), nullptr, 10);
  if (bins > H2OK_MAX_BINS) {
    fprintf(stderr, "error: Invalid value for '--bins <bins>': Error: at most %d bins.\n", H2OK_MAX_BINS);
    return 1;
  }

  std::vector<int32_t> devs(gpus);
  for (int i = 0; i < gpus; i++) devs[i] = i;
  h2ok_group* g = nullptr;
  OK(h2ok_hist_group_create(devs.data(), gpus, &g));
  OK(h2ok_hist_group_load_synth(g, shape_id, seed, rows, (int32_t)cols, 0));
  h2ok_params p;
  memset(&p, 0, sizeof(p));
  p.nbins = p.nbins_top_level = (int32_t)bins;
  p.max_depth = (int32_t)depth;
  p.min_rows = 10.0;
  p.min_split_improvement = 1e-5;
  OK(h2ok_hist_group_loop_create(g, &p, nullptr));
  OK(h2ok_hist_group_set_options(g, H2OK_REDUCE_SCATTER, 1, 0));

  double sec = 0, hist_ms = 0;
  long long levels = 0;
  uint64_t sums[3] = {0, 0, 0};
  for (int t = 0; t <= trees; t++) {  // tree 0 warms up
    OK(h2ok_hist_group_begin_tree_synth(g, seed, 0));
    OK(h2ok_hist_group_sync(g));
    const auto t0 = std::chrono::steady_clock::now();
    int32_t nch = 1, lv = 0;
    double hm = 0;
    while (nch > 0 && lv < depth) {
      OK(h2ok_hist_group_level(g, nullptr, nullptr, &nch));
      float ms = 0;
      OK(h2ok_hist_group_hist_ms(g, &ms));
      hm += ms;
      lv++;
    }
    OK(h2ok_hist_group_sync(g));
    const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (t > 0) {
      sec += s;
      hist_ms += hm;
      levels += lv;
    }
    if (t == trees) OK(h2ok_hist_group_checksum(g, sums));
  }
  OK(h2ok_hist_group_destroy(g));

  const double ms_level = sec * 1e3 / (double)levels;
  const double cells_s = (double)rows * (double)cols * (double)levels / sec;
  if (isatty(STDOUT_FILENO)) {
    printf("GBM histogram path, %lld rows x %lld cols, %lld bins, depth %lld, %d GPU(s)\n", rows, cols, bins, depth, gpus);
    printf("  %-34s %12.3f\n", "ms per tree level", ms_level);
    printf("  %-34s %12.3f\n", "of which histogram pass (ms)", hist_ms / (double)levels);
    printf("  %-34s %12.1f\n", "G cells / s per tree level", cells_s / 1e9);
    printf("  %-34s %016llx %016llx %016llx\n", "checksum (acc, splits, leaf)", (unsigned long long)sums[0],
           (unsigned long long)sums[1], (unsigned long long)sums[2]);
  } else {
    printf("{\"metric\": \"gbm_histogram_rows_x_cols_per_s_per_tree_level\", \"value\": %.6e, \"unit\": \"cells/s\", "
           "\"n_gpus\": %d, \"ms_per_level\": %.6f, \"hist_ms_per_level\": %.6f, \"levels\": %lld, \"rows\": %lld, "
           "\"cols\": %lld, \"bins\": %lld, \"depth\": %lld, \"checksum\": [\"%016llx\", \"%016llx\", \"%016llx\"]}\n",
           cells_s, gpus, ms_level, hist_ms / (double)levels, levels, rows, cols, bins, depth, (unsigned long long)sums[0],
           (unsigned long long)sums[1], (unsigned long long)sums[2]);
  }
  return 0;
}
