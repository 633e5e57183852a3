// rtrun — `rtrun -n N ./scene [args]`: starts N ranks of a DistributedMain scene on one box, one per GPU,
// the way `mpirun -n N` does for the reference (scenes/spheres.cu:83-98 binds rank r to device r % count).
// Each child gets RT_RANK / RT_WORLD_SIZE (read by the mpi.h stand-in) and a shared RT_JOB_ID for the NCCL
// rendezvous file.  Exit status: the first non-zero child status, else 0.
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

#include <ctime>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

int main(int argc, char **argv) {
  int n = 1, arg = 1;
  if (argc >= 3 && (!strcmp(argv[1], "-n") || !strcmp(argv[1], "-np"))) {
    n = atoi(argv[2]);
    arg = 3;
  }
  if (n < 1 || arg >= argc) {
    fprintf(stderr, "usage: rtrun -n N program [args...]\n");
    return 2;
  }
  // per-run nonce of the rendezvous file: pid + wall-clock start in ns (a reused pid has another start time)
  struct timespec ts;
  clock_gettime(CLOCK_REALTIME, &ts);
  std::string job = std::to_string((long)getpid()) + "-" + std::to_string((long long)ts.tv_sec * 1000000000LL + ts.tv_nsec);
  std::vector<pid_t> pids;
  for (int r = 0; r < n; r++) {
    pid_t pid = fork();
    if (pid < 0) {
      perror("fork");
      return 1;
    }
    if (pid == 0) {
      setenv("RT_RANK", std::to_string(r).c_str(), 1);
      setenv("RT_WORLD_SIZE", std::to_string(n).c_str(), 1);
      setenv("RT_JOB_ID", job.c_str(), 1);
      execvp(argv[arg], argv + arg);
      perror("execvp");
      _exit(127);
    }
    pids.push_back(pid);
  }
  int rc = 0;
  for (pid_t pid : pids) {
    int st = 0;
    if (waitpid(pid, &st, 0) < 0) continue;
    int code = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + (WIFSIGNALED(st) ? WTERMSIG(st) : 0);
    if (code != 0 && rc == 0) rc = code;
  }
  return rc;
}
