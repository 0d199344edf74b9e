// oracle/shim: hand-written equivalent of the header the reference's CMake
// generates from ext/kmerind/src/config/config.hpp.in (no MPI, no OpenMP).
#ifndef CONFIG_H
#define CONFIG_H
#define PROJ_SRC_DIR "/root/reference"
#define PROJ_BIN_DIR "/root/repo/oracle/_ref"
#define OMP_SHARE_DEFAULT
#define BLISS_UNUSED(x) do { (void)(x); } while(0)
#define BLISS_ALIGNED_ARRAY(name, count, alignsize) name alignas(alignsize) [count]
#define BLISS_ALIGNED_VAR(name, alignsize) name alignas(alignsize)
#endif
