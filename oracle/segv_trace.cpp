// Test infrastructure: prints a backtrace on SIGSEGV/SIGABRT/SIGBUS (the GPU box has no debugger); linked
// into the reference test programs built by `make reftests`.  Installed by an early-priority constructor
// and run on an alternate stack so that a stack overflow and a crash in another static constructor are
// reported too.
#include <execinfo.h>
#include <initializer_list>
#include <signal.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>
namespace {
void on_fatal(int sig, siginfo_t* info, void*)
{
    void* frames[64];
    const int n = backtrace(frames, 64);
    char msg[96] = "\n*** fatal signal at address 0x";
    size_t len = strlen(msg);
    unsigned long a = (unsigned long)info->si_addr;
    for (int s = 60; s >= 0; s -= 4) msg[len++] = "0123456789abcdef"[(a >> s) & 15];
    memcpy(msg + len, ", backtrace:\n", 13);
    (void)!write(2, msg, len + 13);
    backtrace_symbols_fd(frames, n, 2);
    signal(sig, SIG_DFL);
    raise(sig);
}
__attribute__((constructor(101))) void install()
{
    static char alt[1 << 16];
    stack_t ss{};
    ss.ss_sp = alt;
    ss.ss_size = sizeof alt;
    sigaltstack(&ss, nullptr);
    struct sigaction sa{};
    sa.sa_sigaction = on_fatal;
    sa.sa_flags = SA_SIGINFO | SA_ONSTACK;
    sigemptyset(&sa.sa_mask);
    for (int s : {SIGSEGV, SIGABRT, SIGBUS, SIGILL, SIGFPE}) sigaction(s, &sa, nullptr);
    void* warm[4];
    backtrace(warm, 4);          // loads libgcc's unwinder now, not inside the handler
}
}  // namespace
