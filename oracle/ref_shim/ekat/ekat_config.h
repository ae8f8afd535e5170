// Stand-in for the CMake-generated ekat_config.h (externals/ekat/src/ekat/ekat_config.h.in): serial host build,
// no MPI, no GPU, no YAML, BFB defaults on (EKAT_DEFAULT_BFB is what the reference's own CPU test builds set).
#ifndef EKAT_CONFIG_H
#define EKAT_CONFIG_H
#define EKAT_DEFAULT_BFB
#endif
