// gqp_common.h - glue of the measurement-only probe library (profiles/probe/libgqprobe.so): peak rates of this GPU
// that bench.py uses as roofline denominators and prototypes that are timed before they are adopted.  Nothing
// here is linked into gpqhe_b200/libgpqhe_b200.so or reachable from include/gpqhe_b200.h.
#pragma once
#include <string>

#define GQ_NO_PDL 1      // the probe kernels do not carry the dependent-launch calls
#include "gq_arith.h"     // gpqhe_b200/csrc: the same arithmetic helpers as the product kernels

#if defined(__GNUC__)
#define GQP_EXPORT __attribute__((visibility("default")))
#else
#define GQP_EXPORT
#endif

void gqp_set_error(const std::string &msg);
extern "C" GQP_EXPORT const char *gqp_last_error(void);
