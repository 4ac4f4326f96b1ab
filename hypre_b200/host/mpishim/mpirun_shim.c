/* hypre_b200/host/mpishim/mpirun_shim.c -- launcher for the message-passing shim in this directory.
 *
 *    mpirun_shim -np N program [args...]
 *
 * creates one connected Unix stream socket pair per pair of ranks, forks N copies of the
 * program and tells each its rank, the world size and the descriptor that leads to every
 * peer (SHIM_RANK, SHIM_SIZE, SHIM_FDS).  Exit status: 0 if every rank exited 0, else the
 * first non-zero status (the other ranks are terminated).
 */
#define _GNU_SOURCE
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

int main(int argc, char **argv)
{
   int np = 1, a = 1;
   while (a < argc && argv[a][0] == '-')
   {
      if ((!strcmp(argv[a], "-np") || !strcmp(argv[a], "-n")) && a + 1 < argc) { np = atoi(argv[a + 1]); a += 2; }
      else { fprintf(stderr, "mpirun_shim: unknown option %s\n", argv[a]); return 2; }
   }
   if (a >= argc || np < 1 || np > 64) { fprintf(stderr, "usage: mpirun_shim -np N program [args...]\n"); return 2; }

   int (*fd)[64] = calloc(64, sizeof(*fd));          /* fd[i][j]: descriptor rank i uses to talk to rank j */
   for (int i = 0; i < np; i++)
   {
      fd[i][i] = -1;
      for (int j = i + 1; j < np; j++)
      {
         int sv[2];
         if (socketpair(AF_UNIX, SOCK_STREAM, 0, sv) != 0) { perror("socketpair"); return 2; }
         fd[i][j] = sv[0]; fd[j][i] = sv[1];
      }
   }
   pid_t pid[64];
   for (int r = 0; r < np; r++)
   {
      pid[r] = fork();
      if (pid[r] < 0) { perror("fork"); return 2; }
      if (pid[r] == 0)
      {
         char buf[64], fds[64 * 12] = "";
         for (int i = 0; i < np; i++)
         {
            for (int j = 0; j < np; j++) { if (i != r && i != j && fd[i][j] >= 0) { close(fd[i][j]); } }
         }
         for (int j = 0; j < np; j++) { snprintf(buf, sizeof(buf), "%s%d", j ? "," : "", fd[r][j]); strcat(fds, buf); }
         snprintf(buf, sizeof(buf), "%d", r);  setenv("SHIM_RANK", buf, 1);
         snprintf(buf, sizeof(buf), "%d", np); setenv("SHIM_SIZE", buf, 1);
         setenv("SHIM_FDS", fds, 1);
         execvp(argv[a], &argv[a]);
         perror("execvp");
         _exit(127);
      }
   }
   for (int i = 0; i < np; i++) { for (int j = 0; j < np; j++) { if (i != j) { close(fd[i][j]); } } }
   int result = 0, left = np;
   while (left > 0)
   {
      int st = 0;
      pid_t p = wait(&st);
      if (p < 0) { break; }
      left--;
      int code = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + (WIFSIGNALED(st) ? WTERMSIG(st) : 0);
      if (code != 0 && result == 0)
      {
         result = code;
         for (int r = 0; r < np; r++) { if (pid[r] != p) { kill(pid[r], SIGTERM); } }
      }
   }
   return result;
}
