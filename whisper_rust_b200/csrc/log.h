#pragma once
namespace wb {
typedef void (*wb_log_callback)(int level, const char * text, void * user_data);
void log_set(wb_log_callback cb, void * ud);
void log_msg(int level, const char * fmt, ...);
void log_error(const char * fmt, ...);
}  // namespace wb
