/* launch.c -- comd-launch: start N ranks of a program on this node (the image has no mpirun).
 *   comd-launch -n N [--] program args...
 * Each child gets COMD_RANK / COMD_NRANKS / COMD_LOCAL_RANK / COMD_JOB (and the torchrun-style
 * RANK / WORLD_SIZE / LOCAL_RANK); the launcher's exit code is the first non-zero child status.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/wait.h>
#include <unistd.h>

int main(int argc, char** argv)
{
   int n = 1, first = 1;
   if (argc > 2 && !strcmp(argv[1], "-n")) { n = atoi(argv[2]); first = 3; }
   if (first < argc && !strcmp(argv[first], "--")) ++first;
   if (n < 1 || first >= argc) { fprintf(stderr, "usage: comd-launch -n N [--] program args...\n"); return 2; }
   char job[64];
   snprintf(job, sizeof job, "%d", (int)getpid());
   for (int r = 0; r < n; ++r)
   {
      pid_t pid = fork();
      if (pid < 0) { perror("fork"); return 1; }
      if (pid == 0)
      {
         char v[32];
         snprintf(v, sizeof v, "%d", r);
         setenv("COMD_RANK", v, 1); setenv("RANK", v, 1); setenv("COMD_LOCAL_RANK", v, 1); setenv("LOCAL_RANK", v, 1);
         snprintf(v, sizeof v, "%d", n);
         setenv("COMD_NRANKS", v, 1); setenv("WORLD_SIZE", v, 1);
         setenv("COMD_JOB", job, 1);
         execvp(argv[first], argv + first);
         perror(argv[first]);
         _exit(127);
      }
   }
   int rc = 0, status;
   while (wait(&status) > 0)
   {
      const int c = WIFEXITED(status) ? WEXITSTATUS(status) : 128 + WTERMSIG(status);
      if (c && !rc) rc = c;
   }
   return rc;
}
