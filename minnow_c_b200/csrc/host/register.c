/* register.c -- a working algorithm registry with the reference's interface
 * (/root/reference/src/register.h:9-49; the reference's own bodies,
 * src/register.c:21-84, are stubs). Host-side C of libminnow_host.so.
 *
 * A Register is passed by value everywhere and every caller in
 * src/funcs.c:122-189 starts from Register_New(), so the table lives here:
 * Register_New() hands out a view (Cap = 0: not owned) of a process-wide
 * table that starts with the built-in algorithms and grows through
 * Register_Add().
 */
#include "minnow_register.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "minnow_cuda.h"

#define MNW_MAX_ALGOS 16
#define MNW_MAX_VERSIONS 16

#define HOST_PANIC(...)                                                      \
    do {                                                                     \
        fprintf(stderr, "Panic at file %s, line %d:\n", __FILE__, __LINE__); \
        fprintf(stderr, __VA_ARGS__);                                        \
        fprintf(stderr, "\n");                                               \
        exit(1);                                                             \
    } while (0)

static uint32_t g_algo[MNW_MAX_ALGOS];
static uint32_t g_version[MNW_MAX_ALGOS][MNW_MAX_VERSIONS];
static RegisterFuncs g_funcs[MNW_MAX_ALGOS][MNW_MAX_VERSIONS];
static void *g_funcPtr[MNW_MAX_ALGOS][MNW_MAX_VERSIONS];
static U32Seq g_versionSeq[MNW_MAX_ALGOS];
static PtrSeq g_funcSeq[MNW_MAX_ALGOS];
static int32_t g_algoLen = 0;
static int g_init = 0;

static void table_add(uint32_t algo, uint32_t version, RegisterFuncs funcs) {
    int32_t a = 0;
    while (a < g_algoLen && g_algo[a] != algo) a++;
    if (a == g_algoLen) {
        if (g_algoLen == MNW_MAX_ALGOS) HOST_PANIC("Register is full (%d algorithms).", MNW_MAX_ALGOS);
        g_algo[a] = algo;
        g_versionSeq[a].Data = g_version[a];
        g_versionSeq[a].Len = 0;
        g_versionSeq[a].Cap = 0;
        g_funcSeq[a].Data = g_funcPtr[a];
        g_funcSeq[a].Len = 0;
        g_funcSeq[a].Cap = 0;
        g_algoLen++;
    }
    int32_t v = 0;
    while (v < g_versionSeq[a].Len && g_version[a][v] != version) v++;
    if (v == g_versionSeq[a].Len) {
        if (v == MNW_MAX_VERSIONS) HOST_PANIC("Register is full (%d versions of %x).", MNW_MAX_VERSIONS, algo);
        g_versionSeq[a].Len++;
        g_funcSeq[a].Len++;
    }
    g_version[a][v] = version;
    g_funcs[a][v] = funcs;
    g_funcPtr[a][v] = &g_funcs[a][v];
}

static void table_init(void) {
    if (g_init) return;
    g_init = 1;
    RegisterFuncs diff;
    diff.CFunc = MnwCudaCompress_v0_9;
    diff.CAlloc = MnwCudaCAlloc_v0_9;
    diff.CFree = MnwCudaCFree_v0_9;
    diff.DFunc = MnwCudaDecompress_v0_9;
    diff.DAlloc = MnwCudaDAlloc_v0_9;
    diff.DFree = MnwCudaDFree_v0_9;
    table_add(MNWCUDA_ALGO_CODE, MNWCUDA_ALGO_VERSION, diff);
}

static const RegisterFuncs *lookup(Register reg, uint32_t algo, uint32_t version) {
    for (int32_t a = 0; a < reg.Algo.Len; a++) {
        if (reg.Algo.Data[a] != algo) continue;
        for (int32_t v = 0; v < reg.Version.Data[a].Len; v++)
            if (reg.Version.Data[a].Data[v] == version)
                return (const RegisterFuncs *)reg.Funcs.Data[a].Data[v];
    }
    return NULL;
}

Register Register_New(void) {
    table_init();
    Register reg;
    reg.Algo.Data = g_algo;
    reg.Algo.Len = g_algoLen;
    reg.Algo.Cap = 0;
    reg.Version.Data = g_versionSeq;
    reg.Version.Len = g_algoLen;
    reg.Version.Cap = 0;
    reg.Funcs.Data = g_funcSeq;
    reg.Funcs.Len = g_algoLen;
    reg.Funcs.Cap = 0;
    return reg;
}

void Register_Free(Register reg) { (void)reg; /* a view: nothing is owned */ }

/* `reg` is a by-value view, so the addition goes to the process-wide table
 * and is seen by every Register_New() from now on. */
void Register_Add(Register reg, uint32_t algo, uint32_t version, RegisterFuncs funcs) {
    (void)reg;
    table_init();
    table_add(algo, version, funcs);
}

bool Register_Supports(Register reg, uint32_t algo, uint32_t version) {
    return lookup(reg, algo, version) != NULL;
}

/* the largest version word of `algo`, compared like semver_Cmp
 * (src/semver.c:9-15: the stage byte is masked out); 0 if unknown */
uint32_t Register_Newest(Register reg, uint32_t algo) {
    uint32_t best = 0;
    for (int32_t a = 0; a < reg.Algo.Len; a++) {
        if (reg.Algo.Data[a] != algo) continue;
        for (int32_t v = 0; v < reg.Version.Data[a].Len; v++) {
            uint32_t w = reg.Version.Data[a].Data[v];
            if ((w & 0x00ffffffu) >= (best & 0x00ffffffu)) best = w;
        }
    }
    return best;
}

Compressor Register_GetCompressor(Register reg, uint32_t algo, uint32_t version) {
    const RegisterFuncs *f = lookup(reg, algo, version);
    if (!f) HOST_PANIC("Version %x of algorithm %x is not registered.", version, algo);
    Compressor c;
    c.Buffer = f->CAlloc ? f->CAlloc() : NULL;
    c.CFunc = f->CFunc;
    return c;
}

Decompressor Register_GetDecompressor(Register reg, uint32_t algo, uint32_t version) {
    const RegisterFuncs *f = lookup(reg, algo, version);
    if (!f) HOST_PANIC("Version %x of algorithm %x is not registered.", version, algo);
    Decompressor d;
    d.Buffer = f->DAlloc ? f->DAlloc() : NULL;
    d.DFunc = f->DFunc;
    return d;
}

void Register_FreeCompressor(Register reg, uint32_t algo, uint32_t version, Compressor comp) {
    const RegisterFuncs *f = lookup(reg, algo, version);
    if (f && f->CFree) f->CFree(comp.Buffer);
}

void Register_FreeDecompressor(Register reg, uint32_t algo, uint32_t version, Decompressor decomp) {
    const RegisterFuncs *f = lookup(reg, algo, version);
    if (f && f->DFree) f->DFree(decomp.Buffer);
}
