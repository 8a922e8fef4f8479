#include "common.h"
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <sys/stat.h>
#include <sys/select.h>
#include <unistd.h>
#include <cuda_runtime.h>

namespace tclb {

static int g_rank = 0, g_print_level = L_OUTPUT, g_root_level = 6;
static bool g_tty_out = false, g_tty_err = false;
static const auto g_start = std::chrono::steady_clock::now();

void log_init(int rank, int print_level, int only_root_level) {
    g_rank = rank;
    g_print_level = print_level;
    g_root_level = only_root_level;
    g_tty_out = isatty(1);
    g_tty_err = isatty(2);
    if (const char *e = getenv("TCLB_LOG_LEVEL")) g_print_level = atoi(e);
}
void log_set_level(int l) { g_print_level = l; }
int log_rank() { return g_rank; }

void vlogf(int level, bool all_ranks, const char *fmt, va_list ap) {
    if (level < g_print_level) return;
    if (!all_ranks && level < g_root_level && g_rank != 0) return;
    static const char *color[] = {"\033[2m", "\033[2m", "\033[2m", "", "\033[32m", "\033[1;32m", "\033[33m", "\033[1;33m", "\033[31m", "\033[1;31m"};
    static const char *tag[] = {"D0", "D1", "D2", "  ", "  ", "  ", "W ", "W!", "E ", "E!"};
    FILE *f = level >= L_WARNING ? stderr : stdout;
    bool tty = level >= L_WARNING ? g_tty_err : g_tty_out;
    char buf[4096];
    vsnprintf(buf, sizeof(buf), fmt, ap);
    size_t n = strlen(buf);
    bool nl = n > 0 && buf[n - 1] == '\n';
    if (nl) buf[n - 1] = 0;
    if (tty) fprintf(f, "%s[%2d] %s%s\033[0m\n", color[level], g_rank, tag[level], buf);
    else fprintf(f, "[%2d] %s%s\n", g_rank, tag[level], buf);
    fflush(f);
}
void logf(int level, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vlogf(level, false, fmt, ap);
    va_end(ap);
}
void fail(const char *fmt, ...) {
    char buf[4096];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    throw Error(buf);
}
std::string strf(const char *fmt, ...) {
    char buf[8192];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    return buf;
}
double walltime() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - g_start).count();
}
bool key_pressed(int *ch) {
    if (!isatty(0)) return false;
    fd_set fds;
    FD_ZERO(&fds);
    FD_SET(0, &fds);
    timeval tv{0, 0};
    if (select(1, &fds, nullptr, nullptr, &tv) > 0) {
        int c = getchar();
        if (ch) *ch = c;
        return true;
    }
    return false;
}
void cuda_check(int err, const char *what, const char *file, int line) {
    if (err != 0) {
        const char *msg = cudaGetErrorString((cudaError_t)err);
        fail("CUDA error %d (%s) in %s at %s:%d", err, msg, what, file, line);
    }
}
std::vector<std::string> split(const std::string &s, char sep) {
    std::vector<std::string> out;
    size_t a = 0;
    while (true) {
        size_t b = s.find(sep, a);
        if (b == std::string::npos) { out.push_back(s.substr(a)); break; }
        out.push_back(s.substr(a, b - a));
        a = b + 1;
    }
    return out;
}
std::string trim(const std::string &s) {
    size_t a = s.find_first_not_of(" \t\r\n");
    if (a == std::string::npos) return "";
    size_t b = s.find_last_not_of(" \t\r\n");
    return s.substr(a, b - a + 1);
}
bool starts_with(const std::string &s, const std::string &p) { return s.size() >= p.size() && s.compare(0, p.size(), p) == 0; }
bool ends_with(const std::string &s, const std::string &p) { return s.size() >= p.size() && s.compare(s.size() - p.size(), p.size(), p) == 0; }
std::string basename_noext(const std::string &path) {
    size_t a = path.find_last_of('/');
    std::string b = a == std::string::npos ? path : path.substr(a + 1);
    size_t d = b.find_last_of('.');
    return d == std::string::npos ? b : b.substr(0, d);
}
void make_parent_dirs(const std::string &path) {
    for (size_t i = 1; i < path.size(); i++)
        if (path[i] == '/') mkdir(path.substr(0, i).c_str(), 0775);
}
}  // namespace tclb
