// Host-side 2-bit packing of sequence bytes before the host -> device copy ("2-bit pre-pack
// on host to cut H2D 4x").  Internal to libtjb_gpu.so; not part of the C ABI.
//
// The reads of one sample are ~1 byte per base of "ACGT" plus one '\n' per read, and the copy over PCIe (~55 GB/s) is
// what bounds tjb_count_sample_host().  A chunk is therefore sent as
//   packed[j]  = the gray codes (byte >> 1) & 3 = A0 C1 T2 G3 of bytes 4j .. 4j+3, byte 4j in bits 0-1, and
//   exceptions = (position, byte) of EVERY byte that is not one of "ACGT" (separators, N, lower case, anything),
// from which the device rebuilds the chunk byte for byte (unpack_kernel / unpack_exc_kernel in tjb_gpu.cu) before the
// unchanged scan kernel runs on it: results cannot differ from the unpacked path.
#pragma once
#include <stddef.h>
#include <stdint.h>

struct tjb_packer;

tjb_packer *tjb_packer_create (int n_threads);   // n_threads <= 0: one per hardware thread, at most 32
void tjb_packer_destroy (tjb_packer *p);
int tjb_packer_threads (const tjb_packer *p);
// workers spin between begin and end (a wake-up through a condition variable costs more than packing a chunk share)
void tjb_packer_begin (tjb_packer *p);
void tjb_packer_end (tjb_packer *p);
// Packs src[0, n) into packed[0, (n + 3) / 4) and the exceptions into exc_pos / exc_val (capacity cap).  Returns the
// number of exceptions, or -1 when they do not fit (the caller then sends the chunk as it is).
long tjb_packer_run (tjb_packer *p, const uint8_t *src, size_t n, uint8_t *packed, uint32_t *exc_pos, uint8_t *exc_val, size_t cap);
// dst[0, n) = src[0, n) by the pool and the caller together (between begin and end), 64 KiB blocks, streaming stores
void tjb_packer_copy (tjb_packer *p, void *dst, const void *src, size_t n);
// dst[40 i .. 40 i + 40) = src[24 i .. 24 i + 24) followed by four int32 -1, for i < n_rows (the result elements cross PCIe
// without their constant tail); _rows = on the calling thread alone
void tjb_packer_expand (tjb_packer *p, void *dst, const void *src, size_t n_rows);
void tjb_expand_rows (void *dst, const void *src, size_t n_rows);
// single-threaded, portable restatement (tests, small inputs)
long tjb_pack_scalar (const uint8_t *src, size_t n, uint8_t *packed, uint32_t *exc_pos, uint8_t *exc_val, size_t cap, uint32_t pos0);
