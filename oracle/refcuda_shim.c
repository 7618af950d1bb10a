/* oracle/refcuda_shim.c -- TEST INFRASTRUCTURE ONLY.
 * The reference's tcq4_k32.cu object needs exactly one non-CUDA symbol, ggml_abort
 * (ggml/src/ggml.c), reached only from CUDA_CHECK failures.  This shim provides it so the fork's
 * kernels can be linked into oracle/_ref/libtcq4_refcuda.so without building ggml-cuda. */
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>

void ggml_abort(const char *file, int line, const char *fmt, ...) {
    va_list ap;
    fprintf(stderr, "[refcuda] abort at %s:%d: ", file, line);
    va_start(ap, fmt);
    vfprintf(stderr, fmt, ap);
    va_end(ap);
    fputc('\n', stderr);
    abort();
}
