/*
 * ORACLE — TEST INFRASTRUCTURE ONLY (see the header of ref_besteffort.c for the rule:
 * the product never includes, links or calls anything under oracle/).
 *
 * CPU restatement of the replica-id helpers and the replica-aware prioritiser of the
 * reference, cmd/nvidia-device-plugin/replica.go, and of the XID list parser,
 * cmd/nvidia-device-plugin/nvidia.go:274-294.
 *
 * PARITY STATUS: pinned.  tests/test_oracle_replica.py runs these functions against every
 * case of the reference's own tables: Test_prioritizeDevices (15 cases,
 * cmd/nvidia-device-plugin/replica_test.go:37-96), Test_stripReplicas (3 cases, :120-122)
 * and TestGetAdditionalXids (10 cases, cmd/nvidia-device-plugin/nvidia_test.go:27-66),
 * transcribed into tests/golden/ by tests/golden/make_golden.py.
 *
 * Go strings are byte strings and sort.Strings orders them byte-wise, which is what
 * strcmp() does on NUL-free input.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define REF_ERR_NONE 0
#define REF_ERR_NON_UNIQUE 1     /* &NonUniqueError{} — result is still returned (replica.go:193-197) */
#define REF_ERR_NO_DEVICES_LEFT 2 /* errors.New("no devices left to allocate") (replica.go:183) */
#define REF_ERR_MUST_MISSING 3   /* fmt.Errorf("device '%s' in mustIncludeDeviceIDs is missing ...") (:129,:136) */
#define REF_ERR_PANIC 4          /* make([]string, len(must), size) with len > cap (replica.go:121) */

static const char joinStr[] = "-replica-"; /* replica.go:26 */

/* replica.go:28-30 — strings.Split(id, joinStr)[0]: everything before the first joinStr */
static char *stripReplica(const char *deviceReplica)
{
    const char *p = strstr(deviceReplica, joinStr);
    size_t n = p ? (size_t)(p - deviceReplica) : strlen(deviceReplica);
    char *r = (char *)malloc(n + 1);
    memcpy(r, deviceReplica, n);
    r[n] = 0;
    return r;
}

static int cmp_str(const void *a, const void *b)
{
    return strcmp(*(const char *const *)a, *(const char *const *)b);
}

/* pack strings NUL-separated into buf; returns 0 or -1 when it does not fit */
static int pack(char *buf, int buflen, char *const *strs, int n)
{
    int off = 0;
    for (int i = 0; i < n; i++) {
        int l = (int)strlen(strs[i]) + 1;
        if (off + l > buflen)
            return -1;
        memcpy(buf + off, strs[i], (size_t)l);
        off += l;
    }
    return 0;
}

/* replica.go:32-45.  out: NUL-separated strings in outbuf, *n_out of them. */
int ref_strip_replicas(const char *const *ids, int n, char *outbuf, int outbuf_len, int *n_out)
{
    char **deviceIDs = (char **)malloc(sizeof(char *) * (size_t)(n + 1));
    int cnt = 0;
    for (int i = 0; i < n; i++) {
        char *devID = stripReplica(ids[i]);
        int exists = 0;
        for (int j = 0; j < cnt && !exists; j++)
            exists = strcmp(deviceIDs[j], devID) == 0;
        if (!exists)
            deviceIDs[cnt++] = devID;
        else
            free(devID);
    }
    qsort(deviceIDs, (size_t)cnt, sizeof(char *), cmp_str); /* :43 sort.Strings */
    int rc = pack(outbuf, outbuf_len, deviceIDs, cnt);
    *n_out = cnt;
    for (int i = 0; i < cnt; i++)
        free(deviceIDs[i]);
    free(deviceIDs);
    return rc;
}

/* replica.go:62-65 */
typedef struct {
    char *key; /* map key: the raw (physical) device id */
    int Allocated;
    const char **names; /* ReplicaDeviceNames: a Go slice = window [off, off+len) */
    int off, len;
} devCount;

/* replica.go:47-54 */
static int find(const char **a, int n, const char *x)
{
    for (int i = 0; i < n; i++)
        if (strcmp(x, a[i]) == 0)
            return i;
    return n;
}

/* replica.go:68-76 with remove() of :56-60 (swap-with-last) */
static int devCount_allocate(devCount *d, const char *allocatedDevice)
{
    const char **s = d->names + d->off;
    int idx = find(s, d->len, allocatedDevice);
    if (idx == d->len)
        return 0;
    s[idx] = s[d->len - 1];
    d->len--;
    d->Allocated = 1;
    return 1;
}

/* replica.go:79-84 */
static const char *devCount_allocateAny(devCount *d)
{
    const char *allocatedDevice = d->names[d->off];
    d->off++;
    d->len--;
    d->Allocated = 1;
    return allocatedDevice;
}

static int cmp_devcount_key(const void *a, const void *b)
{
    return strcmp(((const devCount *)a)->key, ((const devCount *)b)->key);
}

/*
 * replica.go:96-198.  Returns the error kind; on REF_ERR_NONE / REF_ERR_NON_UNIQUE the
 * sorted result is in outbuf (NUL-separated, *n_out strings); on REF_ERR_MUST_MISSING
 * errbuf holds the offending id (the %s of the message).  Go returns a nil slice with the
 * other errors.
 */
int ref_prioritize_devices(const char *const *availableDeviceIDs, int n_avail, const char *const *mustIncludeDeviceIDs,
                           int n_must, int allocationSize, char *outbuf, int outbuf_len, int *n_out, char *errbuf,
                           int errbuf_len)
{
    *n_out = 0;
    if (errbuf_len > 0)
        errbuf[0] = 0;
    int rc = REF_ERR_NONE;

    /* :98-112 — map[string]*devCount as a flat table (iteration order never matters: :115
     * sorts inside each entry and :143-147 sorts the keys) */
    devCount *raw = (devCount *)calloc((size_t)(n_avail + 1), sizeof(devCount));
    int nraw = 0;
    for (int i = 0; i < n_avail; i++) {
        char *dev = stripReplica(availableDeviceIDs[i]);
        int e = -1;
        for (int j = 0; j < nraw && e < 0; j++)
            if (strcmp(raw[j].key, dev) == 0)
                e = j;
        if (e >= 0) {
            raw[e].names[raw[e].len++] = availableDeviceIDs[i];
            free(dev);
        } else {
            raw[nraw].key = dev;
            raw[nraw].Allocated = 0;
            raw[nraw].names = (const char **)malloc(sizeof(char *) * (size_t)n_avail);
            raw[nraw].off = 0;
            raw[nraw].len = 0;
            raw[nraw].names[raw[nraw].len++] = availableDeviceIDs[i];
            nraw++;
        }
    }
    /* :115-117 */
    for (int j = 0; j < nraw; j++)
        qsort(raw[j].names, (size_t)raw[j].len, sizeof(char *), cmp_str);

    const char **allocated = (const char **)malloc(sizeof(char *) * (size_t)((allocationSize > n_must ? allocationSize : n_must) + 1));
    int n_alloc = 0;

    /* :121 make([]string, len(must), allocationSize) */
    if (allocationSize < 0 || n_must > allocationSize) {
        rc = REF_ERR_PANIC;
        goto done;
    }
    n_alloc = n_must;
    int unique = 1;

    /* :126-139 */
    for (int i = 0; i < n_must; i++) {
        const char *deviceID = mustIncludeDeviceIDs[i];
        char *dev = stripReplica(deviceID);
        devCount *deviceCount = 0;
        for (int j = 0; j < nraw && !deviceCount; j++)
            if (strcmp(raw[j].key, dev) == 0)
                deviceCount = &raw[j];
        free(dev);
        if (!deviceCount) {
            snprintf(errbuf, (size_t)errbuf_len, "%s", deviceID);
            rc = REF_ERR_MUST_MISSING;
            goto done;
        }
        if (deviceCount->Allocated)
            unique = 0;
        if (!devCount_allocate(deviceCount, deviceID)) {
            snprintf(errbuf, (size_t)errbuf_len, "%s", deviceID);
            rc = REF_ERR_MUST_MISSING;
            goto done;
        }
        allocated[i] = deviceID;
    }

    /* :143-147 */
    qsort(raw, (size_t)nraw, sizeof(devCount), cmp_devcount_key);

    /* :149-190 */
    for (int i = n_alloc; i < allocationSize; i++) {
        int allocatedHighest = 0, unallocatedHighest = 0;
        devCount *leastUtilizedDevAllocated = 0, *leastUtilizedDevUnallocated = 0;
        for (int j = 0; j < nraw; j++) {
            devCount *deviceCount = &raw[j];
            int count = deviceCount->len;
            if (deviceCount->Allocated) {
                if (count > allocatedHighest) {
                    leastUtilizedDevAllocated = deviceCount;
                    allocatedHighest = count;
                }
            } else {
                if (count > unallocatedHighest) {
                    leastUtilizedDevUnallocated = deviceCount;
                    unallocatedHighest = count;
                }
            }
        }
        devCount *deviceCount;
        if (leastUtilizedDevUnallocated)
            deviceCount = leastUtilizedDevUnallocated;
        else if (leastUtilizedDevAllocated)
            deviceCount = leastUtilizedDevAllocated;
        else {
            rc = REF_ERR_NO_DEVICES_LEFT;
            goto done;
        }
        if (deviceCount->Allocated)
            unique = 0;
        allocated[n_alloc++] = devCount_allocateAny(deviceCount);
    }
    /* :191 */
    qsort(allocated, (size_t)n_alloc, sizeof(char *), cmp_str);
    if (pack(outbuf, outbuf_len, (char *const *)allocated, n_alloc) != 0) {
        rc = -1;
        goto done;
    }
    *n_out = n_alloc;
    /* :193-197 */
    if (!unique)
        rc = REF_ERR_NON_UNIQUE;

done:
    for (int j = 0; j < nraw; j++) {
        free(raw[j].key);
        free(raw[j].names);
    }
    free(raw);
    free(allocated);
    return rc;
}

/* unicode.IsSpace, which strings.TrimSpace trims by: '\t' '\n' '\v' '\f' '\r' ' ', U+0085, U+00A0, U+1680,
 * U+2000..U+200A, U+2028, U+2029, U+202F, U+205F, U+3000.  Returns the UTF-8 length of the white-space rune that
 * starts at p (n bytes are readable), 0 if there is none. */
static int go_space_len(const unsigned char *p, long n)
{
    if (n <= 0)
        return 0;
    if (*p == ' ' || (*p >= '\t' && *p <= '\r'))
        return 1;
    if (n >= 2 && p[0] == 0xC2 && (p[1] == 0x85 || p[1] == 0xA0))
        return 2;
    if (n >= 3) {
        if (p[0] == 0xE1 && p[1] == 0x9A && p[2] == 0x80)
            return 3;
        if (p[0] == 0xE2 && p[1] == 0x80 && ((p[2] >= 0x80 && p[2] <= 0x8A) || p[2] == 0xA8 || p[2] == 0xA9 || p[2] == 0xAF))
            return 3;
        if (p[0] == 0xE2 && p[1] == 0x81 && p[2] == 0x9F)
            return 3;
        if (p[0] == 0xE3 && p[1] == 0x80 && p[2] == 0x80)
            return 3;
    }
    return 0;
}

/*
 * nvidia.go:274-294.  strconv.ParseUint(s, 10, 64): one or more decimal digits, no sign,
 * no underscores (base is explicit), value <= 2^64-1; anything else is "malformed" and
 * skipped.  Returns the number of xids written to out (order kept).
 */
int ref_get_additional_xids(const char *input, uint64_t *out, int out_cap)
{
    int n = 0;
    if (input[0] == 0) /* :275-277 */
        return 0;
    const char *p = input;
    for (;;) {
        const char *q = strchr(p, ',');
        size_t len = q ? (size_t)(q - p) : strlen(p);
        /* :281 TrimSpace */
        const unsigned char *b = (const unsigned char *)p, *e = b + len;
        int adv;
        while ((adv = go_space_len(b, e - b)) > 0)
            b += adv;
        for (;;) {
            int cut = 0;
            for (int len = 1; len <= 3 && len <= e - b && !cut; len++)
                if (go_space_len(e - len, len) == len)
                    cut = len;
            if (!cut)
                break;
            e -= cut;
        }
        if (b < e) { /* :282-284 */
            int ok = 1;
            uint64_t v = 0;
            for (const unsigned char *c = b; c < e && ok; c++) {
                if (*c < '0' || *c > '9') {
                    ok = 0;
                    break;
                }
                unsigned d = (unsigned)(*c - '0');
                if (v > (UINT64_MAX - d) / 10) /* range error */
                    ok = 0;
                else
                    v = v * 10 + d;
            }
            if (ok && n < out_cap) /* :285-290 */
                out[n++] = v;
        }
        if (!q)
            break;
        p = q + 1;
    }
    return n;
}
