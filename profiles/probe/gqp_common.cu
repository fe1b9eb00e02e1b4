#include "gqp_common.h"
static thread_local std::string g_gqp_err;
void gqp_set_error(const std::string &msg) { g_gqp_err = msg; }
extern "C" GQP_EXPORT const char *gqp_last_error(void) { return g_gqp_err.c_str(); }
