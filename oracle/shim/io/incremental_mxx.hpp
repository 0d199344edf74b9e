// oracle/shim: empty stand-in shadowing ext/kmerind/src/io/incremental_mxx.hpp,
// which include/kmerhash/hashmap_robinhood.hpp:37 includes only for code that is
// commented out there.  Keeps <mpi.h> off the oracle build.  TEST INFRASTRUCTURE ONLY.
#ifndef KH_ORACLE_SHIM_INCREMENTAL_MXX_HPP_
#define KH_ORACLE_SHIM_INCREMENTAL_MXX_HPP_
#endif
