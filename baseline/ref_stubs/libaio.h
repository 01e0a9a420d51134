/* Type-only stand-in for <libaio.h>: the reference's AIOHandler.h is included by its Merger headers; the hybrid
 * (disk) merge is not part of the timed online merge and AIOHandler.cc is not compiled into the reference arm. */
#ifndef UDA_REF_STUB_LIBAIO_H
#define UDA_REF_STUB_LIBAIO_H
#include <stdint.h>
#include <sys/types.h>
#include <time.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef struct io_context* io_context_t;
struct iocb { void* data; unsigned key; short aio_lio_opcode; short aio_reqprio; int aio_fildes; union { struct { void* buf; unsigned long nbytes; long long offset; } c; } u; };
struct io_event { void* data; struct iocb* obj; unsigned long res; unsigned long res2; };
#ifdef __cplusplus
}
#endif
#endif
