// Test infrastructure: prints a backtrace on SIGSEGV/SIGABRT (the GPU box has no debugger); linked into
// the reference test programs built by `make reftests`.
#include <execinfo.h>
#include <signal.h>
#include <unistd.h>
namespace {
void on_fatal(int sig)
{
    void* frames[64];
    const int n = backtrace(frames, 64);
    const char msg[] = "\n*** fatal signal, backtrace:\n";
    (void)!write(2, msg, sizeof msg - 1);
    backtrace_symbols_fd(frames, n, 2);
    signal(sig, SIG_DFL);
    raise(sig);
}
struct Install { Install() { signal(SIGSEGV, on_fatal); signal(SIGABRT, on_fatal); signal(SIGBUS, on_fatal); } } g_install;
}  // namespace
