/* Hand-written config.h for compiling the reference's dav1d C sources in place
 * (x86-64 Linux, C path only: HAVE_ASM=0, both bitdepths). Mirrors what
 * cmake/dav1d.cmake:258-530 would generate; it is host-library build input. */
#pragma once
#define ARCH_AARCH64 0
#define ARCH_ARM 0
#define ARCH_PPC64LE 0
#define ARCH_RISCV 0
#define ARCH_RV32 0
#define ARCH_RV64 0
#define ARCH_LOONGARCH 0
#define ARCH_LOONGARCH32 0
#define ARCH_LOONGARCH64 0
#define ARCH_X86 1
#define ARCH_X86_32 0
#define ARCH_X86_64 1
#define CONFIG_8BPC 1
#define CONFIG_16BPC 1
#define CONFIG_LOG 1
#define CONFIG_MACOS_KPERF 0
#define ENDIANNESS_BIG 0
#define HAVE_ASM 0
#define HAVE_AVX512ICL 0
#define HAVE_C11_GENERIC 1
#define HAVE_CLOCK_GETTIME 1
#define HAVE_DLSYM 1
#define HAVE_POSIX_MEMALIGN 1
#define HAVE_MEMALIGN 1
#define HAVE_ALIGNED_ALLOC 1
#define HAVE_UNISTD_H 1
#define HAVE_SYS_TYPES_H 1
#define HAVE_IO_H 0
#define HAVE_PTHREAD_NP_H 0
#define HAVE_GETAUXVAL 1
#define HAVE_ELF_AUX_INFO 0
#define HAVE_PTHREAD_GETAFFINITY_NP 1
#define HAVE_PTHREAD_SETAFFINITY_NP 1
#define HAVE_PTHREAD_SETNAME_NP 1
#define HAVE_PTHREAD_SET_NAME_NP 0
#define HAVE_SIGACTION 1
#define STACK_ALIGNMENT 16
#define TRIM_DSP_FUNCTIONS 0
