// Logging, error handling and small utilities shared by the runtime core.
// Counterpart of the reference's console layer (src/Global.cpp.Rt:91-246, macros
// src/Global.h.Rt:114-165): 10 levels, rank prefix, root-only below a threshold, colour on a tty.
#pragma once
#include <cstdarg>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <vector>

namespace tclb {

enum LogLevel { L_DEBUG0 = 0, L_DEBUG1, L_DEBUG2, L_OUTPUT, L_NOTICE, L_NOTICE_IMP, L_WARNING, L_WARNING_IMP, L_ERROR, L_FATAL };

void log_init(int rank, int print_level, int only_root_level = 6);
void log_set_level(int print_level);
int log_rank();
void logf(int level, const char *fmt, ...) __attribute__((format(printf, 2, 3)));
void vlogf(int level, bool all_ranks, const char *fmt, va_list ap);

#define TCLB_DEBUG(...) ::tclb::logf(::tclb::L_DEBUG2, __VA_ARGS__)
#define TCLB_OUTPUT(...) ::tclb::logf(::tclb::L_OUTPUT, __VA_ARGS__)
#define TCLB_NOTICE(...) ::tclb::logf(::tclb::L_NOTICE_IMP, __VA_ARGS__)
#define TCLB_WARN(...) ::tclb::logf(::tclb::L_WARNING_IMP, __VA_ARGS__)
#define TCLB_ERROR(...) ::tclb::logf(::tclb::L_ERROR, __VA_ARGS__)

struct Error : public std::runtime_error {
    explicit Error(const std::string &s) : std::runtime_error(s) {}
};
[[noreturn]] void fail(const char *fmt, ...) __attribute__((format(printf, 1, 2)));
std::string strf(const char *fmt, ...) __attribute__((format(printf, 1, 2)));

double walltime();           // seconds since process start (steady clock)
bool key_pressed(int *ch);   // non-blocking stdin poll (reference kbhit)

// CUDA error check helper (only evaluated when a CUDA call is made)
void cuda_check(int err, const char *what, const char *file, int line);
#define TCLB_CUDA(call) ::tclb::cuda_check((int)(call), #call, __FILE__, __LINE__)

std::vector<std::string> split(const std::string &s, char sep);
std::string trim(const std::string &s);
bool starts_with(const std::string &s, const std::string &p);
bool ends_with(const std::string &s, const std::string &p);
std::string basename_noext(const std::string &path);
void make_parent_dirs(const std::string &path);

}  // namespace tclb
