/* distlearn_loader.h — C ABI of the native batch loader in libdistlearn_host.so (csrc/loader.cpp).
 *
 * Counterpart of the reference's `dataset` package as its examples use it (Dataset{partition, partitions} +
 * sampledBatcher{samplerKind, batchSize} -> getBatch(); examples/mnist.lua:26-40, examples/cifar10.lua:41-71).
 * Worker threads assemble batch b — a pure function of (seed, b) — into slot b % depth of caller-owned buffers
 * (page-locked when the batches go to a GPU), converting uint8 samples to float32 with x = u8 * scale[c] + shift[c].
 * All functions return 0 on success, a negative value on error (text in dll_last_error(), per thread). */
#ifndef DISTLEARN_LOADER_H
#define DISTLEARN_LOADER_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* dll_last_error(void);

/* data: [n_total, sample_elems] uint8 (is_u8 = 1) or float32; labels: int64 [n_total] in [0, num_classes).
 * channels divides sample_elems; scale / shift: `channels` floats each (NULL = 1 / 0).
 * partition (1-based) of partitions: this loader serves samples partition-1, partition-1+partitions, ...
 * kind: 0 = permutation (new shuffle per epoch, short last batch), 1 = label-uniform (uniform class, then uniform member).
 * x_slots / y_slots: `depth` (>= 2) buffers of batch * sample_elems floats / batch int64.  The caller keeps data, labels
 * and the slots alive until dll_destroy. */
int dll_create(const void* data, int is_u8, const int64_t* labels, int64_t n_total, int sample_elems, int channels,
               const float* scale, const float* shift, int num_classes, int partition, int partitions, int kind, int batch,
               uint64_t seed, int threads, int depth, float* const* x_slots, int64_t* const* y_slots, void** out);

/* Next batch, strictly in order: which slot holds it, how many samples it has, its number.  Blocks until a worker has
 * finished it; -2 after timeout_s seconds (<= 0: wait for ever); -1 if all `depth` slots are outstanding. */
int dll_next(void* loader, int* slot, int* size, int64_t* batch_no, double timeout_s);

/* Return the OLDEST outstanding batch's slot to the workers (anything still reading it, e.g. an asynchronous H2D copy,
 * must have finished). */
int dll_release(void* loader);

int64_t dll_num_batches(void* loader);      /* per epoch */
int64_t dll_num_samples(void* loader);      /* in this partition */
uint64_t dll_samples_done(void* loader);    /* assembled so far */
int dll_destroy(void* loader);              /* stops and joins the workers */

/* CIFAR-10 binary batch file (records of 1 label byte + 3072 pixel bytes) -> data[n, 3072], labels[n];
 * returns the number of records read (<= max_records) or -1. */
int64_t dll_read_cifar_bin(const char* path, uint8_t* data, int64_t* labels, int64_t max_records);

#ifdef __cplusplus
}
#endif
#endif
