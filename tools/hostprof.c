/* hostprof.c — SIGPROF sampling profile of the host side (parse + entropy decode + command
 * recording) with the NULL emulation engine; development aid, not part of the product.
 *   gcc -O2 -g -o /tmp/hostprof tools/hostprof.c -ldl
 *   EAV1_EMU_NULL=1 /tmp/hostprof tests/emu/libdav1d_b200_emu.so tests/golden/c2_1080p_bench.ivf 5 | sort | uniq -c | sort -rn | head -40
 * prints one "lib+offset" line per sample; resolve with addr2line -f -e lib offset. */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <signal.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>
#include <time.h>
#include <ucontext.h>

#define MAXS 200000
static void *samples[MAXS], *callers[MAXS];
static volatile int n_samples;
static void on_prof(int sig, siginfo_t *si, void *uc_)
{
    ucontext_t *uc = uc_;
    if (n_samples < MAXS) {
        /* the word at the stack pointer: the return address when the sample sits in a frameless leaf (memcpy, memset) */
        callers[n_samples] = *(void **)uc->uc_mcontext.gregs[REG_RSP];
        samples[n_samples++] = (void *)uc->uc_mcontext.gregs[REG_RIP];
    }
}
typedef struct { int w, h, layout, bpc, frame_type, show_frame, showable; int64_t offset; } Info;
int main(int argc, char **argv)
{
    void *lib = dlopen(argv[1], RTLD_NOW | RTLD_GLOBAL);
    if (!lib) { fprintf(stderr, "%s\n", dlerror()); return 1; }
    void *(*open_)(const uint8_t *, size_t, int, int, int) = dlsym(lib, "eav1_stream_open");
    int (*next_)(void *, Info *, uint8_t *, size_t, size_t *) = dlsym(lib, "eav1_stream_next");
    void (*close_)(void *) = dlsym(lib, "eav1_stream_close");
    FILE *f = fopen(argv[2], "rb"); fseek(f, 0, SEEK_END); long sz = ftell(f); rewind(f);
    uint8_t *ivf = malloc(sz), *obu = malloc(sz); fread(ivf, 1, sz, f); fclose(f);
    size_t pos = ivf[6] | ivf[7] << 8, n = 0;
    while (pos + 12 <= (size_t)sz) { uint32_t s; memcpy(&s, ivf + pos, 4); memcpy(obu + n, ivf + pos + 12, s); n += s; pos += 12 + s; }
    int reps = argc > 3 ? atoi(argv[3]) : 3;
    size_t cap = 1920 * 1088 * 4; uint8_t *buf = malloc(cap);
    struct sigaction sa = {0}; sa.sa_sigaction = on_prof; sa.sa_flags = SA_SIGINFO | SA_RESTART; sigaction(SIGPROF, &sa, NULL);
    struct itimerval it = {{0, 500}, {0, 500}}; setitimer(ITIMER_PROF, &it, NULL);
    struct timespec t0, t1; clock_gettime(CLOCK_MONOTONIC, &t0);
    int frames = 0;
    for (int r = 0; r < reps; r++) {
        void *h = open_(obu, n, 1, 7, 1);
        if (!h) { fprintf(stderr, "open failed\n"); return 1; }
        Info info; size_t used;
        while (next_(h, &info, buf, cap, &used) == 0) frames++;
        close_(h);
    }
    clock_gettime(CLOCK_MONOTONIC, &t1);
    struct itimerval off = {{0, 0}, {0, 0}}; setitimer(ITIMER_PROF, &off, NULL);
    double dt = (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec);
    fprintf(stderr, "%d frames in %.3f s = %.1f fps, %d samples\n", frames, dt, frames / dt, n_samples);
    for (int i = 0; i < n_samples; i++) {
        Dl_info di;
        if (dladdr(samples[i], &di) && di.dli_fname && strstr(di.dli_fname, "libc.so") && dladdr(callers[i], &di) &&
            di.dli_fname && !strstr(di.dli_fname, "libc.so")) {
            const char *b = strrchr(di.dli_fname, '/'); b = b ? b + 1 : di.dli_fname;
            printf("libc<-%s +0x%lx\n", b, (unsigned long)((char *)callers[i] - (char *)di.dli_fbase));
            continue;
        }
        if (dladdr(samples[i], &di) && di.dli_fname) {
            const char *b = strrchr(di.dli_fname, '/'); b = b ? b + 1 : di.dli_fname;
            printf("%s +0x%lx\n", b, (unsigned long)((char *)samples[i] - (char *)di.dli_fbase));
        } else printf("? %p\n", samples[i]);
    }
    return 0;
}
