// bp_internal.h — shared between the translation units of libbp_b200.so (not part of the C ABI).
#ifndef BP_INTERNAL_H
#define BP_INTERNAL_H
// Records the calling thread's error message (what bp_last_error() returns) and returns `code`.
int bp_fail(int code, const char *fmt, ...);
#endif
