/*
 * papi.h -- empty stand-in for the three PAPI calls in the reference's
 * performance/Measurements.cpp (no PAPI on this image).  Counters read 0.
 * TEST INFRASTRUCTURE, see oracle/README.md.
 */
#ifndef ORACLE_STUB_PAPI_H_
#define ORACLE_STUB_PAPI_H_
typedef long long long_long;
#define PAPI_OK 0
#define PAPI_TOT_CYC 1
#define PAPI_L1_TCM 2
#define PAPI_L1_DCM 3
static inline int PAPI_start_counters(int *events, int n) { (void)events; (void)n; return PAPI_OK; }
static inline int PAPI_stop_counters(long_long *values, int n) { for (int i = 0; i < n; ++i) values[i] = 0; return PAPI_OK; }
static inline const char *PAPI_strerror(int e) { (void)e; return "papi stub"; }
#endif
