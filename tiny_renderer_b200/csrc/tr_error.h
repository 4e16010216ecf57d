// tr_error.h — thread-local error message behind tr_last_error() (include/tinyr.h).
#pragma once
#include <cstdarg>

namespace tr {
// Stores a printf-formatted message for the calling thread and returns `status` (so callers can
// `return fail(TR_ERR_IO, "...")`).
int fail(int status, const char* fmt, ...);
void clear_error();
const char* last_error();
}  // namespace tr
