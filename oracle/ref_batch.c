/*
 * ORACLE — TEST INFRASTRUCTURE ONLY (see ref_besteffort.c).  Batch driver: runs the oracle
 * over an array of packed container requests (the 32-byte gpa_request records of
 * include/gpualloc.h), sequentially or sharded over host threads.  It exists so that
 * tests and bench.py's cpu_baseline / --impl reference legs can time and compare the CPU
 * restatement of the reference's loop over r.ContainerRequests
 * (cmd/nvidia-device-plugin/server.go:273-311 -> Policy.Allocate at :298) without Python
 * call overhead.  The Go reference handles one request at a time on one goroutine; the
 * threaded mode shards INDEPENDENT requests, which is the most parallelism the reference's
 * design admits.
 */
#include <pthread.h>
#include <stdint.h>
#include <string.h>

int ref_besteffort_fast(int nt, const int32_t *W, const uint8_t *nlinks, const int32_t *avail, int n_avail,
                        const int32_t *req, int n_req, int size, int mode, int32_t *out, int *n_out,
                        unsigned long long *visited);

typedef struct {
    uint8_t avail[16];
    uint16_t req_mask;
    uint8_t n_avail, size, flags, n_req_extra;
    uint8_t reserved[10];
} packed_request;

typedef struct {
    int8_t dev[16];
} packed_result;

#define RES_END (-128)
#define RES_ERROR (-127)
#define REQ_FOREIGN 1

static void run_one(int nt, const int32_t *W, const uint8_t *nlinks, const packed_request *r, int mode, packed_result *o,
                    unsigned long long *visited)
{
    int32_t avail[16], req[40], out[64];
    int n_req = 0, n_out = 0;
    int n = r->n_avail > 16 ? 16 : r->n_avail;
    for (int i = 0; i < n; i++)
        avail[i] = r->avail[i];
    for (int p = 0; p < n; p++)
        if (r->req_mask >> p & 1)
            req[n_req++] = avail[p];
    if (r->flags & REQ_FOREIGN) {
        int extra = r->n_req_extra ? r->n_req_extra : 1;
        for (int i = 0; i < extra && n_req < 40; i++)
            req[n_req++] = -2; /* REF_FOREIGN */
    }
    unsigned long long v = 0;
    int rc = ref_besteffort_fast(nt, W, nlinks, avail, n, req, n_req, r->size, mode, out, &n_out, &v);
    if (visited)
        *visited += v;
    memset(o->dev, RES_END, sizeof o->dev);
    if (rc != 0) {
        o->dev[0] = RES_ERROR;
        o->dev[1] = (int8_t)(rc == 1 ? -4 /* GPA_ERR_ASYMMETRIC */ : -1);
        return;
    }
    for (int i = 0; i < n_out && i < 16; i++)
        o->dev[i] = (int8_t)out[i];
}

typedef struct {
    int nt;
    const int32_t *W;
    const uint8_t *nlinks;
    const packed_request *reqs;
    packed_result *out;
    long long begin, end;
    int mode;
    unsigned long long visited;
} shard_t;

static void *shard_main(void *p)
{
    shard_t *s = (shard_t *)p;
    for (long long i = s->begin; i < s->end; i++)
        run_one(s->nt, s->W, s->nlinks, &s->reqs[i], s->mode, &s->out[i], &s->visited);
    return 0;
}

/* threads <= 1: sequential on the calling thread.  Returns candidate partitions visited. */
unsigned long long ref_besteffort_batch(int nt, const int32_t *W, const uint8_t *nlinks, const void *reqs, long long count,
                                        int mode, void *out, int threads)
{
    if (threads < 1)
        threads = 1;
    if (threads > 256)
        threads = 256;
    shard_t sh[256];
    pthread_t th[256];
    long long per = (count + threads - 1) / threads;
    for (int t = 0; t < threads; t++) {
        sh[t].nt = nt;
        sh[t].W = W;
        sh[t].nlinks = nlinks;
        sh[t].reqs = (const packed_request *)reqs;
        sh[t].out = (packed_result *)out;
        sh[t].begin = per * t < count ? per * t : count;
        sh[t].end = per * (t + 1) < count ? per * (t + 1) : count;
        sh[t].mode = mode;
        sh[t].visited = 0;
    }
    if (threads == 1) {
        shard_main(&sh[0]);
        return sh[0].visited;
    }
    for (int t = 0; t < threads; t++)
        pthread_create(&th[t], 0, shard_main, &sh[t]);
    unsigned long long v = 0;
    for (int t = 0; t < threads; t++) {
        pthread_join(th[t], 0);
        v += sh[t].visited;
    }
    return v;
}
