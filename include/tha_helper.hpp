// tha_helper.hpp -- error-check macros of the host layer (reference: include/hip_helper.hpp:4-11).
// CHECK_HIP wrapped hip* runtime calls; here the runtime is reached through the C-ABI
// (tha_b200.h), whose calls return a thablasStatus_t value, so the macro checks that.
#pragma once
#include <stdio.h>
#include <stdlib.h>

#include "tha_b200.h"

#define CHECK_THA(cmd)                                                                                      \
  do {                                                                                                      \
    int status_ = (cmd);                                                                                    \
    if (status_ != THA_STATUS_SUCCESS) {                                                                    \
      fprintf(stderr, "THA Error: %s (%d): %s:%d\n", tha_last_error(), status_, __FILE__, __LINE__);       \
      fflush(stdout);                                                                                       \
      exit(EXIT_FAILURE);                                                                                   \
    }                                                                                                       \
  } while (0)

// Source compatibility for callers written against the reference.
#define CHECK_HIP(cmd) CHECK_THA(cmd)

#define MAX_NUM_SUPPORTED_GPUS 32
#define WARP_SIZE 32          /* the reference's 64 is the gfx908 wavefront */
#define MAX_BLOCK_SIZE 1024
