// ORACLE BUILD SHIM (test infrastructure): LOG(severity) << ... to stderr.
#ifndef ORACLE_SHIM_ABSL_LOG_H_
#define ORACLE_SHIM_ABSL_LOG_H_
#include <iostream>
#define LOG(sev) std::cerr
#endif
