// Stand-in for the CMake-generated scream_config.h (components/eamxx/src/scream_config.h.in).
// Pack sizes 1: results do not depend on the pack size, and size 1 is what the GPU build uses.
#ifndef SCREAM_CONFIG_H
#define SCREAM_CONFIG_H
#ifndef P3REF_SINGLE_PRECISION
#define SCREAM_DOUBLE_PRECISION
#endif
#ifndef SCREAM_PACK_SIZE
#define SCREAM_PACK_SIZE 1
#endif
#ifndef SCREAM_SMALL_PACK_SIZE
#define SCREAM_SMALL_PACK_SIZE SCREAM_PACK_SIZE
#endif
#define SCREAM_POSSIBLY_NO_PACK_SIZE 1
#define SCREAM_NUM_VERTICAL_LEV 72
#define SCREAM_TEST_LEVEL 0
#define SCREAM_MPI_ON_DEVICE 0
#ifndef SCREAM_DATA_DIR
#define SCREAM_DATA_DIR "."
#endif
#define EAMXX_GIT_VERSION "ref"
#define EAMXX_VERSION "0.0.0"
#endif
