/* shim: the reference includes this header (apps/kmeans/kmeans.cpp:3) but never calls it */
#ifndef MINIPS_SHIM_PROFILER_H_
#define MINIPS_SHIM_PROFILER_H_
static inline int ProfilerStart(const char*) { return 0; }
static inline void ProfilerStop(void) {}
#endif
