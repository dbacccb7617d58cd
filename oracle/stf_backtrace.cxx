// oracle/stf_backtrace.cxx -- TEST INFRASTRUCTURE: linked into oracle/_ref/stf_gpu only.  When the process dies on an uncaught
// exception or a fatal signal, print the call stack (function names through -rdynamic) so that a failure of the file-level
// parity test on the GPU box, where there is no debugger, says where it happened.
#include <csignal>
#include <cstdio>
#include <cstdlib>
#include <exception>
#include <execinfo.h>
#include <unistd.h>

namespace {
void dump(const char *why)
{
    void *frames[64];
    const int n = backtrace(frames, 64);
    fprintf(stderr, "stf_gpu: %s; call stack:\n", why);
    backtrace_symbols_fd(frames, n, STDERR_FILENO);
}
void on_terminate()
{
    dump("uncaught exception");
    abort();
}
void on_signal(int sig)
{
    dump(sig == SIGSEGV ? "SIGSEGV" : "fatal signal");
    signal(sig, SIG_DFL);
    raise(sig);
}
struct Install {
    Install()
    {
        std::set_terminate(on_terminate);
        signal(SIGSEGV, on_signal);
        signal(SIGFPE, on_signal);
    }
} install;
} // namespace
