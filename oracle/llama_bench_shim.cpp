// oracle/llama_bench_shim.cpp -- TEST INFRASTRUCTURE ONLY.
// The reference's tools/llama-bench/llama-bench.cpp (compiled unmodified, in place) calls five helpers of the reference's
// `common` library (common/common.h:78,606,625,627 and the build-info symbols).  Building all of common/ would pull chat templates,
// JSON and curl into the oracle; these are our own minimal definitions of exactly those helpers, so that llama-bench links against
// libllama-ref.so alone.  Behaviour relevant to llama-bench: thread-count default, "%"-formatting, whitespace strip, hex CPU masks.
#include "common.h"

#include <cstdarg>
#include <cstdio>
#include <thread>

int          LLAMA_BUILD_NUMBER  = 0;
const char * LLAMA_COMMIT        = "oracle";
const char * LLAMA_COMPILER      = "g++";
const char * LLAMA_BUILD_TARGET  = "x86_64-linux-gnu";

int32_t cpu_get_num_math() {
    const unsigned n = std::thread::hardware_concurrency();
    return n > 0 ? (int32_t) n : 4;
}

std::string string_format(const char * fmt, ...) {
    va_list a, b;
    va_start(a, fmt);
    va_copy(b, a);
    const int n = vsnprintf(nullptr, 0, fmt, a);
    std::string out(n > 0 ? (size_t) n : 0, '\0');
    if (n > 0) vsnprintf(&out[0], (size_t) n + 1, fmt, b);
    va_end(b);
    va_end(a);
    return out;
}

std::string string_strip(const std::string & s) {
    size_t b = 0, e = s.size();
    while (b < e && isspace((unsigned char) s[b])) ++b;
    while (e > b && isspace((unsigned char) s[e - 1])) --e;
    return s.substr(b, e - b);
}

bool parse_cpu_mask(const std::string & mask, bool (&boolmask)[GGML_MAX_N_THREADS]) {
    size_t start = (mask.size() > 1 && mask[0] == '0' && (mask[1] == 'x' || mask[1] == 'X')) ? 2 : 0;
    const size_t ndig = mask.size() - start;
    if (ndig == 0 || ndig * 4 > GGML_MAX_N_THREADS) return false;
    for (size_t i = 0; i < ndig; ++i) {
        const char c = mask[start + i];
        int v = c >= '0' && c <= '9' ? c - '0' : c >= 'a' && c <= 'f' ? c - 'a' + 10 : c >= 'A' && c <= 'F' ? c - 'A' + 10 : -1;
        if (v < 0) return false;
        const size_t lo = (ndig - 1 - i) * 4;                      // the last digit holds CPUs 0-3
        for (int b = 0; b < 4; ++b) boolmask[lo + b] = boolmask[lo + b] || ((v >> b) & 1);
    }
    return true;
}

#include <sys/resource.h>
// llama-bench --prio: nice value per priority class (common/common.h: bool set_process_priority(enum ggml_sched_priority))
bool set_process_priority(enum ggml_sched_priority prio) {
    int nice_value = 0;
    switch (prio) {
        case GGML_SCHED_PRIO_LOW:      nice_value = 5;   break;
        case GGML_SCHED_PRIO_NORMAL:   return true;
        case GGML_SCHED_PRIO_MEDIUM:   nice_value = -5;  break;
        case GGML_SCHED_PRIO_HIGH:     nice_value = -10; break;
        case GGML_SCHED_PRIO_REALTIME: nice_value = -20; break;
    }
    return setpriority(PRIO_PROCESS, 0, nice_value) == 0;
}
