// unpack77.h -- host-side 77-bit payload -> text (stand-in for WSJT-X packjt77.f90 unpack77,
// reached in the reference through f_interop.cpp:61-71).
#pragma once
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif
// c77: 77 characters '0'/'1'; msg: 37 characters, blank padded, NOT NUL terminated; *ok = 1/0.
void msk144_unpack77(const char c77[77], const int* nrx, char msg[37], int* ok);
// Forget all hashed call signs (the tables otherwise persist for the life of the process).
void msk144_unpack77_reset(void);
// Per-stream call-sign hash tables: msk144_unpack77_in() is msk144_unpack77() against the tables of `state`
// (NULL = the process-wide set).  The built-in decoder only; an external library keeps its own module state.
void* msk144_hash_state_new(void);
void msk144_hash_state_free(void* state);
void msk144_hash_state_clear(void* state);
void msk144_unpack77_in(void* state, const char c77[77], const int* nrx, char msg[37], int* ok);
// Route text decoding through a user-supplied library exporting __packjt77_MOD_unpack77 (NULL / "" = built-in).
// Returns 0, or -1 if the library or the symbol cannot be loaded (the built-in stays active).
int msk144_set_unpack77_library(const char* path);
int msk144_unpack77_is_external(void);
// gfortran-ABI alias so that code written against WSJT-X links unchanged.
void __packjt77_MOD_unpack77(char c77[], const int* nrx, char msg[], int* ok, size_t len77, size_t len37);
#ifdef __cplusplus
}
#endif
