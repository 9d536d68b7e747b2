/* TEST INFRASTRUCTURE: stand-in for <valgrind/callgrind.h>, which the reference's basic_bench.h includes
 * (src/basic_bench.h:17) for three client-request macros around its timed loop. Outside valgrind they do nothing. */
#pragma once
#define CALLGRIND_START_INSTRUMENTATION
#define CALLGRIND_STOP_INSTRUMENTATION
#define CALLGRIND_TOGGLE_COLLECT
