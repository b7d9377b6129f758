/* main.c -- the `FEM` command line on top of the B200 path.
 *
 * Same sub-commands, options, output file and statistics lines as the reference's
 * src/FEM.c:23-51, src/FEM_index.c:11-39 and src/FEM_map.c:57-227:
 *   FEM index <window_size> <step_size> <reference> <output> [--gpu[=N]]
 *   FEM map -e INT -t INT [-a INT] [-f g] --ref FILE --index FILE --read1 FILE -o FILE
 * Differences, all additive: -t is the number of host threads that build SAM text (the mapping
 * itself runs on the GPUs); --gpus N / FEM_GPUS selects how many devices to use (default 1);
 * the seeding window/step come from the index header unless --asis (or FEM_COMPAT=asis) asks for
 * the unmodified reference's hard-coded 12/3.
 */
#include <getopt.h>
#include <inttypes.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/resource.h>
#include <sys/time.h>

#include "femhost.h"

#define FEM_VERSION "0.2"

static double real_time(void) {
  struct timeval tv;
  gettimeofday(&tv, NULL);
  return (double)tv.tv_sec + 1e-6 * (double)tv.tv_usec;
}
static double cpu_time(void) {
  struct rusage r;
  getrusage(RUSAGE_SELF, &r);
  return (double)(r.ru_utime.tv_sec + r.ru_stime.tv_sec) + 1e-6 * (double)(r.ru_utime.tv_usec + r.ru_stime.tv_usec);
}

static void usage_main(void) {
  fprintf(stderr, "\nProgram: FEM (Fast and Efficient short read Mapper), B200 build\n");
  fprintf(stderr, "Version: %s\n\n", FEM_VERSION);
  fprintf(stderr, "Usage:   FEM <command> [options]\n\n");
  fprintf(stderr, "Command: index   build index for reference\n");
  fprintf(stderr, "         map     map reads\n\n");
  fprintf(stderr, "Note: To use FEM, you need to first index the genome with `FEM index'.\n\n");
}

static void usage_map(void) {
  fprintf(stderr, "\nUsage:  FEM map [options] \n\nOptions:\n");
  fprintf(stderr, "        -e       INT  error threshold \n");
  fprintf(stderr, "        -t       INT  number of threads \n");
  fprintf(stderr, "        -f       STR  seeding algorithm: \"g\" for group seeding and \"v\" for variable-length seeding \n");
  fprintf(stderr, "        -a       INT  # additional q-grams (only for test)\n");
  fprintf(stderr, "        --gpus   INT  number of GPUs (default 1, or $FEM_GPUS)\n");
  fprintf(stderr, "        --asis        seed with window 12 / step 3 whatever the index header says\n");
  fprintf(stderr, "\nInput/output: \n");
  fprintf(stderr, "        --ref    STR  Input reference file\n");
  fprintf(stderr, "        --index  STR  Input index file\n");
  fprintf(stderr, "        --read1  STR  Input read1 file\n");
  fprintf(stderr, "        -o       STR  Output SAM file \n\n");
}

static int index_main(int argc, char **argv) {
  if (argc < 5) {
    fprintf(stderr, "Too few args!\nUsage: FEM index <window_size> <step_size> <reference> <output>\n");
    exit(EXIT_FAILURE);
  }
  const int k = atoi(argv[1]), step = atoi(argv[2]);
  fprintf(stderr, "k: %d, step size: %d, reference: %s, output: %s\n", k, step, argv[3], argv[4]);
  femhost_records ref;
  femhost_records_init(&ref, 0);
  double t0 = real_time();
  if (femhost_load_all(argv[3], &ref)) { fprintf(stderr, "%s\n", femhost_last_error()); exit(EXIT_FAILURE); }
  fprintf(stderr, "Number of sequences: %u\nNumber of bases: %" PRIu64 "\nLoaded all sequences successfully in %fs\n",
          ref.n_records, ref.n_bases, real_time() - t0);
  /* additive: a fifth argument "--gpu[=N]" or FEM_INDEX_GPU=N builds the same file on device N */
  int device = getenv("FEM_INDEX_GPU") ? atoi(getenv("FEM_INDEX_GPU")) : -1;
  if (argc > 5 && strncmp(argv[5], "--gpu", 5) == 0) device = argv[5][5] == '=' ? atoi(argv[5] + 6) : 0;
  if (device >= 0) {
    double sec[3];
    if (femhost_index_build_device(&ref, k, step, device, argv[4], sec)) { fprintf(stderr, "%s\n", femhost_last_error()); exit(EXIT_FAILURE); }
    fprintf(stderr, "Reference resident on GPU %d in %fs.\nBuilt index in %fs.\nSaved index successfully in %fs\n", device, sec[0], sec[1], sec[2]);
    femhost_records_free(&ref);
    return 0;
  }
  t0 = real_time();
  femhost_index idx;
  if (femhost_index_build(&ref, k, step, &idx)) { fprintf(stderr, "%s\n", femhost_last_error()); exit(EXIT_FAILURE); }
  fprintf(stderr, "Lookup table size: %zu, occurrence table size: %zu.\nBuilt index in %fs.\n", idx.lut_n, idx.n_occ, real_time() - t0);
  t0 = real_time();
  if (femhost_index_save(&idx, argv[4])) { fprintf(stderr, "%s\n", femhost_last_error()); exit(EXIT_FAILURE); }
  fprintf(stderr, "Saved index successfully in %fs\n", real_time() - t0);
  femhost_index_free(&idx);
  femhost_records_free(&ref);
  return 0;
}

static int map_main(int argc, char **argv) {
  const char *ref_path = NULL, *index_path = NULL, *read1_path = NULL, *out_path = NULL;
  femhost_map_options opt;
  memset(&opt, 0, sizeof opt);
  opt.params.k = 0; opt.params.step = 0; /* from the index header */
  opt.params.e = 2; opt.params.a = 1;    /* src/FEM_map.c:69-70 */
  opt.n_threads = 1;
  int n_gpus = getenv("FEM_GPUS") ? atoi(getenv("FEM_GPUS")) : 1;
  int asis = getenv("FEM_COMPAT") && strcmp(getenv("FEM_COMPAT"), "asis") == 0;
  static const struct option long_opt[] = {{"help", no_argument, NULL, 'h'},       {"ref", required_argument, NULL, 'r'},
                                           {"index", required_argument, NULL, 'i'}, {"read1", required_argument, NULL, 'b'},
                                           {"gpus", required_argument, NULL, 'G'},  {"asis", no_argument, NULL, 'A'},
                                           {"batch", required_argument, NULL, 'B'}, {NULL, 0, NULL, 0}};
  int c;
  while ((c = getopt_long(argc, argv, "ha:f:e:t:o:r:i:b:", long_opt, NULL)) >= 0) {
    switch (c) {
      case 'r': ref_path = optarg; break;
      case 'i': index_path = optarg; break;
      case 'b': read1_path = optarg; break;
      case 'e': opt.params.e = atoi(optarg); break;
      case 't': opt.n_threads = atoi(optarg); break;
      case 'a': opt.params.a = atoi(optarg); break;
      case 'G': n_gpus = atoi(optarg); break;
      case 'A': asis = 1; break;
      case 'B': opt.batch_reads = (uint32_t)atoi(optarg); break;
      case 'f':
        if (strcmp(optarg, "v") != 0 && strcmp(optarg, "g") != 0) { /* src/FEM_map.c:108-118 */
          fprintf(stderr, "Wrong name of seeding algorithm!\n");
          usage_map();
          exit(EXIT_FAILURE);
        }
        break;
      case 'o': out_path = optarg; break;
      default: usage_map(); exit(EXIT_SUCCESS);
    }
  }
  if (asis) { opt.params.k = 12; opt.params.step = 3; } /* src/FEM_map.c:67-68 */
  /* src/FEM_map.c:29-55 */
  const char *bad = NULL;
  if (opt.params.e < 0 || opt.params.e > 7) bad = "Wrong error threshold.";
  else if (opt.n_threads <= 0) bad = "Wrong number of threads.";
  else if (opt.params.a < 0 || opt.params.a > 2) bad = "Wrong number of additional q-grams.";
  else if (!ref_path) bad = "Reference file path is required.";
  else if (!index_path) bad = "Index file path is required.";
  else if (!read1_path) bad = "Read file path is required.";
  else if (!out_path) bad = "Output file path is required.";
  else if (n_gpus < 1) bad = "Wrong number of GPUs.";
  if (bad) { fprintf(stderr, "%s\n", bad); usage_map(); exit(EXIT_FAILURE); }

  double t0 = real_time();
  femhost_mapper *m = femhost_mapper_create(ref_path, index_path, n_gpus, NULL);
  if (!m) { fprintf(stderr, "%s\n", femhost_last_error()); exit(EXIT_FAILURE); }
  const femhost_records *ref = femhost_mapper_reference(m);
  fprintf(stderr, "Number of sequences: %u\nNumber of bases: %" PRIu64 "\n", ref->n_records, ref->n_bases);
  fprintf(stderr, "Loaded reference and index onto %d GPU(s) in %fs\n", n_gpus, real_time() - t0);

  femgpu_stats st;
  femhost_map_timing tm;
  if (femhost_map_file(m, &opt, read1_path, out_path, &st, &tm)) {
    fprintf(stderr, "%s\n", femhost_last_error());
    femhost_mapper_destroy(m);
    exit(EXIT_FAILURE);
  }
  if (st.num_infeasible_reads)
    fprintf(stderr, "Warning: %" PRIu64 " reads are too short for -e %d with this index (the reference crashes on them); they were left unmapped.\n",
            st.num_infeasible_reads, opt.params.e);
  /* src/FEM_map.c:214-219 */
  fprintf(stderr, "The number of read: %" PRIu64 "\n", st.num_reads);
  fprintf(stderr, "The number of mapped read: %" PRIu64 "\n", st.num_mapped_reads);
  fprintf(stderr, "The number of candidate before additional q-gram filter: %" PRIu64 "\n", st.num_candidates_before_filter);
  fprintf(stderr, "The number of candidate: %" PRIu64 "\n", st.num_candidates);
  fprintf(stderr, "The number of mapping: %" PRIu64 "\n", st.num_mappings);
  fprintf(stderr, "Time: %fs\n", tm.wall);
  fprintf(stderr, "[phases] parse %.3fs, gpu threads %.3fs, SAM text %.3fs over %d threads, write %.3fs; %u batches, %u kernel launches, %" PRIu64 " reads formatted on the device\n",
          tm.parse, tm.gpu_wait, tm.format, opt.n_threads, tm.write, tm.n_batches, tm.n_launches, tm.device_sam_reads);
  femhost_mapper_destroy(m);
  return 0;
}

int main(int argc, char **argv) {
  if (argc < 2) {
    fprintf(stderr, "Too few arguements.\n");
    usage_main();
    exit(EXIT_FAILURE);
  }
  const double t0 = real_time(), c0 = cpu_time();
  int rc;
  if (strcmp(argv[1], "index") == 0) rc = index_main(argc - 1, argv + 1);
  else if (strcmp(argv[1], "map") == 0) rc = map_main(argc - 1, argv + 1);
  else {
    fprintf(stderr, "[main] unrecognized command '%s'\n", argv[1]);
    exit(EXIT_FAILURE);
  }
  if (rc == 0) { /* src/FEM.c:43-48 */
    fprintf(stderr, "[main] Version: %s\n[main] CMD:", FEM_VERSION);
    for (int i = 0; i < argc; ++i) fprintf(stderr, " %s", argv[i]);
    fprintf(stderr, "\n[main] Real time: %.3f sec; CPU: %.3f sec\n", real_time() - t0, cpu_time() - c0);
  }
  return rc;
}
