// oracle/shim: MPI-free stand-in that shadows the reference's
// ext/kmerind/src/utils/benchmark_utils.hpp (which pulls <mpi.h> through
// utils/timer.hpp -> mxx/reduction.hpp).  The table headers only use the
// BL_BENCH_* phase-timer macros from it; here they expand to nothing, exactly
// what the reference itself does when BL_BENCHMARK != 1
// (ext/kmerind/src/utils/benchmark_utils.hpp:50-66).
// TEST INFRASTRUCTURE ONLY - used to compile oracle/_ref from the reference
// sources where they lie under /root/reference.  No reference file is edited.
#ifndef KH_ORACLE_SHIM_BENCHMARK_UTILS_HPP_
#define KH_ORACLE_SHIM_BENCHMARK_UTILS_HPP_
#include <cmath>
#include <iomanip>
#include <iostream>
#define BL_BENCH_INIT(title)
#define BL_BENCH_RESET(title)
#define BL_BENCH_LOOP_START(title, id)
#define BL_BENCH_LOOP_RESUME(title, id)
#define BL_BENCH_LOOP_PAUSE(title, id)
#define BL_BENCH_LOOP_END(title, id, name, n_elem)
#define BL_BENCH_START(title)
#define BL_BENCH_COLLECTIVE_START(title, name, comm)
#define BL_BENCH_COLLECTIVE_END(title, name, n_elem, comm)
#define BL_BENCH_END(title, name, n_elem)
#define BL_BENCH_REPORT(title, rank)
#define BL_BENCH_REPORT_MPI(title, rank, comm)
#define BL_BENCH_REPORT_NAMED(title, name)
#define BL_BENCH_REPORT_MPI_NAMED(title, name, comm)
#endif
