// oracle/shim: hand-written equivalent of the header generated from
// ext/kmerind/src/config/logger_config.hpp.in (no logging, no benchmarking).
#ifndef LOGGER_CONFIG_H
#define LOGGER_CONFIG_H
#define USE_LOGGER BLISS_LOGGING_NO_LOG
#define LOGGER_VERBOSITY BLISS_LOGGER_VERBOSITY_WARNING
#define BL_BENCHMARK 0
#define BL_BENCHMARK_MEM 0
#define BL_BENCHMARK_TIME 0
#endif
