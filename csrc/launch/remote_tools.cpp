// kungfu-rrun / kungfu-distribute: SSH fan-out helpers.
//
//   kungfu-distribute -H ip:slots,... [-u user] <command...>     run a command on every host
//   kungfu-rrun -H ip:slots,... -np N [-u user] [kungfu-run flags] <prog> [args]
//                                                   start `kungfu-run` on every host of the job
// Parity: /root/reference/srcs/go/cmd/kungfu-rrun/rrun.go,
// /root/reference/srcs/go/cmd/kungfu-distribute/kungfu-distribute.go:76-88,
// /root/reference/srcs/go/utils/runner/remote/remote.go:22-131, /root/reference/srcs/go/utils/ssh/ssh.go.
// The reference embeds an SSH client; here the system `ssh` binary is driven
// (BatchMode, no tty), one child per host, output prefixed per host.
#include <fcntl.h>
#include <signal.h>
#include <sys/wait.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "core/plan.hpp"

using namespace kf;

static std::string shell_quote(const std::string &s)
{
    std::string o = "'";
    for (char c : s) {
        if (c == '\'') o += "'\\''";
        else o += c;
    }
    return o + "'";
}

// children of this process, one ssh per host, each in its own process group; when one of them fails the
// others are cancelled (the reference cancels the remaining ssh sessions the same way,
// /root/reference/srcs/go/utils/runner/remote/remote.go:22-57)
static std::mutex g_mu;
static std::vector<pid_t> g_children;
static std::atomic<bool> g_cancelled{false};

static void cancel_others(pid_t except)
{
    if (g_cancelled.exchange(true)) return;
    std::vector<pid_t> victims;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        for (pid_t c : g_children)
            if (c > 0 && c != except) victims.push_back(c);
    }
    for (pid_t c : victims) kill(-c, SIGTERM);
    std::thread([victims] {
        std::this_thread::sleep_for(std::chrono::seconds(3));
        for (pid_t c : victims) kill(-c, SIGKILL);  // gone already: ESRCH, harmless
    }).detach();
}

static int run_on(size_t slot, const std::string &user, const HostSpec &h, const std::string &cmd, bool dry)
{
    const std::string target = (user.empty() ? "" : user + "@") + h.public_addr;
    if (dry) {
        printf("ssh -o BatchMode=yes %s %s\n", target.c_str(), shell_quote(cmd).c_str());
        return 0;
    }
    int pfd[2];
    // close-on-exec: the hosts are started from concurrent threads, and a write end that leaked into a sibling's
    // ssh would keep this host's pipe open (no EOF, so its exit would go unnoticed) until that sibling finished
    if (pipe2(pfd, O_CLOEXEC) != 0) return 1;
    pid_t pid = fork();
    if (pid == 0) {
        setpgid(0, 0);
        dup2(pfd[1], 1);
        dup2(pfd[1], 2);
        close(pfd[0]);
        execlp("ssh", "ssh", "-o", "BatchMode=yes", "-o", "StrictHostKeyChecking=no", target.c_str(), cmd.c_str(), (char *)nullptr);
        fprintf(stderr, "cannot exec ssh: %s\n", strerror(errno));
        _exit(127);
    }
    setpgid(pid, pid);
    {
        std::lock_guard<std::mutex> lk(g_mu);
        g_children[slot] = pid;
    }
    if (g_cancelled) kill(-pid, SIGTERM);  // another host failed while this one was being started
    close(pfd[1]);
    FILE *in   = fdopen(pfd[0], "r");
    char *line = nullptr;
    size_t cap = 0;
    while (getline(&line, &cap, in) > 0) printf("[%s] %s", h.public_addr.c_str(), line);
    free(line);
    fclose(in);
    int st = 0;
    waitpid(pid, &st, 0);
    {
        std::lock_guard<std::mutex> lk(g_mu);
        g_children[slot] = -1;
    }
    const int rc = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + (WIFSIGNALED(st) ? WTERMSIG(st) : 0);
    if (rc != 0 && !g_cancelled) {
        fprintf(stderr, "%s exited with %d: stopping the other hosts\n", h.public_addr.c_str(), rc);
        cancel_others(pid);
    }
    return rc;
}

int main(int argc, char **argv)
{
    setvbuf(stdout, nullptr, _IOLBF, 0);  // the hosts' lines reach a pipe or a log file as they are printed
    const bool rrun = strstr(argv[0], "rrun") != nullptr;
    std::string hosts = "127.0.0.1:1", user, self_flag;
    bool dry = false;
    std::vector<std::string> pass;  // flags forwarded to kungfu-run
    int i = 1;
    for (; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "-H" && i + 1 < argc) hosts = argv[++i];
        else if (a == "-u" && i + 1 < argc) user = argv[++i];
        else if (a == "-dry-run") dry = true;
        else if (rrun && a.size() > 1 && a[0] == '-') {
            pass.push_back(a);
            if (a.find('=') == std::string::npos && a != "-w" && a != "-k" && a != "-q" && a != "-allow-nvlink" && i + 1 < argc)
                pass.push_back(argv[++i]);
        } else break;
    }
    if (i >= argc) {
        fprintf(stderr, "usage: %s -H ip:slots[:public],... [-u user] [-dry-run] %s\n", argv[0],
                rrun ? "[kungfu-run flags] <prog> [args]" : "<command...>");
        return 2;
    }
    HostList hl;
    if (!parse_host_list(hosts, &hl)) {
        fprintf(stderr, "bad -H %s\n", hosts.c_str());
        return 2;
    }
    std::string tail;
    for (; i < argc; ++i) tail += std::string(tail.empty() ? "" : " ") + shell_quote(argv[i]);
    std::vector<int> rc(hl.size(), 0);
    g_children.assign(hl.size(), -1);
    std::vector<std::thread> ths;
    for (size_t k = 0; k < hl.size(); ++k) {
        std::string cmd = tail;
        if (rrun) {
            cmd = "kungfu-run -H " + shell_quote(hosts) + " -self " + format_ipv4(hl[k].ipv4);
            for (auto &f : pass) cmd += " " + shell_quote(f);
            cmd += " " + tail;
        }
        ths.emplace_back([&, k, cmd] { rc[k] = run_on(k, user, hl[k], cmd, dry); });
    }
    for (auto &t : ths) t.join();
    int bad = 0;
    for (size_t k = 0; k < rc.size(); ++k) {
        if (rc[k]) {
            fprintf(stderr, "%s exited with %d\n", hl[k].public_addr.c_str(), rc[k]);
            // hosts that were cancelled end with 128 + SIGTERM: report the failure that caused it
            if (bad == 0 || bad == 128 + SIGTERM || bad == 128 + SIGKILL) bad = rc[k];
        }
    }
    return bad;
}
