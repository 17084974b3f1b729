// Process-global log sink of the library (whisper.cpp keeps the same: whisper.cpp:984-990, 8935-8969).
#include "log.h"

#include <stdarg.h>
#include <stdio.h>

namespace wb {

static wb_log_callback g_cb = nullptr;
static void * g_ud = nullptr;

void log_set(wb_log_callback cb, void * ud) {
    g_cb = cb;
    g_ud = ud;
}

static void vlog(int level, const char * fmt, va_list ap) {
    char buf[2048];
    vsnprintf(buf, sizeof(buf), fmt, ap);
    if (g_cb) g_cb(level, buf, g_ud);
    else {
        fputs(buf, stderr);
        fflush(stderr);
    }
}

void log_msg(int level, const char * fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vlog(level, fmt, ap);
    va_end(ap);
}

void log_error(const char * fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vlog(4 /*GGML_LOG_LEVEL_ERROR*/, fmt, ap);
    va_end(ap);
}

}  // namespace wb
