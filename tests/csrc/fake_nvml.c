/*
 * TEST STAND-IN for libnvidia-ml.so.1: exports the NVML symbols the product dlopen()s
 * (k8s-gpu-sharing-plugin_b200/csrc/nvml_host.cpp) over a scripted topology and an
 * injectable event queue, so that the topology builder and the XID health watcher can be
 * exercised on machines without a driver.  Not part of the product.
 */
#include <pthread.h>
#include <stdio.h>
#include <string.h>
#include <time.h>

#include <nvml.h>

#define MAXD 64
#define MAXL 18

static int g_n;
static char g_uuid[MAXD][96];
static char g_bus[MAXD][32];
static int g_level[MAXD][MAXD];
static int g_link_on[MAXD][MAXL];
static char g_link_bus[MAXD][MAXL][32];
static int g_link_type[MAXD][MAXL];
static int g_links_supported = MAXL; /* links >= this answer NVML_ERROR_INVALID_ARGUMENT */
static int g_events_supported[MAXD];
static unsigned long long g_calls;

typedef struct {
    int dev; /* -1: no device attached */
    unsigned long long etype, data;
} ev_t;
static ev_t g_q[4096];
static int g_qh, g_qt;
static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t g_cv = PTHREAD_COND_INITIALIZER;

/* memory and minor numbers: scripted per device (defaults: 183359 MiB free, 0 used, minor = index) */
static unsigned long long g_mem_free[64], g_mem_used[64];
static int g_minor[64], g_mem_scripted[64], g_minor_scripted[64];

static nvmlDevice_t H(int i) { return (nvmlDevice_t)(size_t)(i + 1); }
static int I(nvmlDevice_t d) { return (int)(size_t)d - 1; }

/* ---- control surface for the tests -------------------------------------------------- */
void fake_nvml_reset(int n)
{
    pthread_mutex_lock(&g_mu);
    g_n = n;
    memset(g_level, 0, sizeof g_level);
    memset(g_link_on, 0, sizeof g_link_on);
    memset(g_link_bus, 0, sizeof g_link_bus);
    g_links_supported = MAXL;
    g_calls = 0;
    g_qh = g_qt = 0;
    memset(g_mem_scripted, 0, sizeof g_mem_scripted);
    memset(g_minor_scripted, 0, sizeof g_minor_scripted);
    for (int i = 0; i < MAXD; i++) {
        snprintf(g_uuid[i], sizeof g_uuid[i], "GPU-%08x-0000-0000-0000-000000000000", i);
        snprintf(g_bus[i], sizeof g_bus[i], "00000000:%02X:00.0", 0x10 + i);
        g_events_supported[i] = 1;
        for (int l = 0; l < MAXL; l++)
            g_link_type[i][l] = NVML_NVLINK_DEVICE_TYPE_GPU;
    }
    pthread_mutex_unlock(&g_mu);
}
void fake_nvml_set_device(int i, const char *uuid, const char *bus)
{
    snprintf(g_uuid[i], sizeof g_uuid[i], "%s", uuid);
    if (bus)
        snprintf(g_bus[i], sizeof g_bus[i], "%s", bus);
}
void fake_nvml_set_level(int i, int j, int level) { g_level[i][j] = level; }
void fake_nvml_set_link(int i, int link, int enabled, const char *remote_bus, int remote_type)
{
    g_link_on[i][link] = enabled;
    snprintf(g_link_bus[i][link], sizeof g_link_bus[i][link], "%s", remote_bus ? remote_bus : "");
    g_link_type[i][link] = remote_type;
}
void fake_nvml_set_links_supported(int n) { g_links_supported = n; }
void fake_nvml_set_events_supported(int i, int yes) { g_events_supported[i] = yes; }
const char *fake_nvml_bus(int i) { return g_bus[i]; }
unsigned long long fake_nvml_calls(void) { return g_calls; }
void fake_nvml_inject(int dev, unsigned long long etype, unsigned long long data)
{
    pthread_mutex_lock(&g_mu);
    g_q[g_qt % 4096].dev = dev;
    g_q[g_qt % 4096].etype = etype;
    g_q[g_qt % 4096].data = data;
    g_qt++;
    pthread_cond_broadcast(&g_cv);
    pthread_mutex_unlock(&g_mu);
}
int fake_nvml_pending(void) { return g_qt - g_qh; }

/* ---- NVML surface --------------------------------------------------------------------- */
nvmlReturn_t nvmlInit_v2(void) { g_calls++; return NVML_SUCCESS; }
nvmlReturn_t nvmlShutdown(void) { return NVML_SUCCESS; }
const char *nvmlErrorString(nvmlReturn_t r)
{
    switch (r) {
    case NVML_SUCCESS: return "Success";
    case NVML_ERROR_NOT_SUPPORTED: return "Not Supported";
    case NVML_ERROR_INVALID_ARGUMENT: return "Invalid Argument";
    case NVML_ERROR_TIMEOUT: return "Timeout";
    case NVML_ERROR_NOT_FOUND: return "Not Found";
    default: return "Unknown Error";
    }
}
nvmlReturn_t nvmlDeviceGetCount_v2(unsigned int *n) { g_calls++; *n = (unsigned)g_n; return NVML_SUCCESS; }
nvmlReturn_t nvmlDeviceGetHandleByIndex_v2(unsigned int i, nvmlDevice_t *d)
{
    g_calls++;
    if ((int)i >= g_n) return NVML_ERROR_INVALID_ARGUMENT;
    *d = H((int)i);
    return NVML_SUCCESS;
}
nvmlReturn_t nvmlDeviceGetHandleByUUID(const char *uuid, nvmlDevice_t *d)
{
    g_calls++;
    for (int i = 0; i < g_n; i++)
        if (strcmp(uuid, g_uuid[i]) == 0) {
            *d = H(i);
            return NVML_SUCCESS;
        }
    return NVML_ERROR_NOT_FOUND;
}
nvmlReturn_t nvmlDeviceGetUUID(nvmlDevice_t d, char *buf, unsigned int len)
{
    g_calls++;
    int i = I(d);
    if (i < 0 || i >= g_n) return NVML_ERROR_INVALID_ARGUMENT;
    snprintf(buf, len, "%s", g_uuid[i]);
    return NVML_SUCCESS;
}
nvmlReturn_t nvmlDeviceGetPciInfo_v3(nvmlDevice_t d, nvmlPciInfo_t *pci)
{
    g_calls++;
    int i = I(d);
    if (i < 0 || i >= g_n) return NVML_ERROR_INVALID_ARGUMENT;
    memset(pci, 0, sizeof *pci);
    snprintf(pci->busId, sizeof pci->busId, "%s", g_bus[i]);
    return NVML_SUCCESS;
}
void fake_nvml_set_memory(int i, unsigned long long free_bytes, unsigned long long used_bytes)
{
    g_mem_free[i] = free_bytes; g_mem_used[i] = used_bytes; g_mem_scripted[i] = 1;
}
void fake_nvml_set_minor(int i, int minor) { g_minor[i] = minor; g_minor_scripted[i] = 1; }
nvmlReturn_t nvmlDeviceGetMemoryInfo(nvmlDevice_t d, nvmlMemory_t *mem)
{
    g_calls++;
    int i = (int)(size_t)d - 1;
    if (i < 0 || i >= g_n) return NVML_ERROR_INVALID_ARGUMENT;
    mem->free = g_mem_scripted[i] ? g_mem_free[i] : 183359ull * 1024 * 1024;
    mem->used = g_mem_scripted[i] ? g_mem_used[i] : 0;
    mem->total = mem->free + mem->used;
    return NVML_SUCCESS;
}
nvmlReturn_t nvmlDeviceGetMinorNumber(nvmlDevice_t d, unsigned int *minor)
{
    g_calls++;
    int i = (int)(size_t)d - 1;
    if (i < 0 || i >= g_n) return NVML_ERROR_INVALID_ARGUMENT;
    if (g_minor_scripted[i] && g_minor[i] < 0) return NVML_ERROR_NOT_SUPPORTED;
    *minor = (unsigned)(g_minor_scripted[i] ? g_minor[i] : i);
    return NVML_SUCCESS;
}

nvmlReturn_t nvmlDeviceGetTopologyCommonAncestor(nvmlDevice_t a, nvmlDevice_t b, nvmlGpuTopologyLevel_t *lvl)
{
    g_calls++;
    int i = I(a), j = I(b);
    if (i < 0 || j < 0 || i >= g_n || j >= g_n) return NVML_ERROR_INVALID_ARGUMENT;
    if (g_level[i][j] < 0) return NVML_ERROR_NOT_SUPPORTED;
    *lvl = (nvmlGpuTopologyLevel_t)g_level[i][j];
    return NVML_SUCCESS;
}
nvmlReturn_t nvmlDeviceGetNvLinkState(nvmlDevice_t d, unsigned int link, nvmlEnableState_t *st)
{
    g_calls++;
    int i = I(d);
    if (i < 0 || i >= g_n || (int)link >= g_links_supported) return NVML_ERROR_INVALID_ARGUMENT;
    *st = g_link_on[i][link] ? NVML_FEATURE_ENABLED : NVML_FEATURE_DISABLED;
    return NVML_SUCCESS;
}
nvmlReturn_t nvmlDeviceGetNvLinkRemotePciInfo_v2(nvmlDevice_t d, unsigned int link, nvmlPciInfo_t *pci)
{
    g_calls++;
    int i = I(d);
    if (i < 0 || i >= g_n || (int)link >= g_links_supported) return NVML_ERROR_INVALID_ARGUMENT;
    memset(pci, 0, sizeof *pci);
    snprintf(pci->busId, sizeof pci->busId, "%s", g_link_bus[i][link]);
    return NVML_SUCCESS;
}
nvmlReturn_t nvmlDeviceGetNvLinkRemoteDeviceType(nvmlDevice_t d, unsigned int link, nvmlIntNvLinkDeviceType_t *t)
{
    g_calls++;
    int i = I(d);
    if (i < 0 || i >= g_n || (int)link >= g_links_supported) return NVML_ERROR_INVALID_ARGUMENT;
    *t = (nvmlIntNvLinkDeviceType_t)g_link_type[i][link];
    return NVML_SUCCESS;
}
nvmlReturn_t nvmlEventSetCreate(nvmlEventSet_t *set) { g_calls++; *set = (nvmlEventSet_t)(size_t)0x5e7; return NVML_SUCCESS; }
nvmlReturn_t nvmlEventSetFree(nvmlEventSet_t set) { (void)set; return NVML_SUCCESS; }
nvmlReturn_t nvmlDeviceRegisterEvents(nvmlDevice_t d, unsigned long long types, nvmlEventSet_t set)
{
    (void)types; (void)set;
    g_calls++;
    int i = I(d);
    if (i < 0 || i >= g_n) return NVML_ERROR_INVALID_ARGUMENT;
    return g_events_supported[i] ? NVML_SUCCESS : NVML_ERROR_NOT_SUPPORTED;
}
nvmlReturn_t nvmlEventSetWait_v2(nvmlEventSet_t set, nvmlEventData_t *data, unsigned int timeoutms)
{
    (void)set;
    struct timespec ts;
    clock_gettime(CLOCK_REALTIME, &ts);
    ts.tv_sec += timeoutms / 1000;
    ts.tv_nsec += (long)(timeoutms % 1000) * 1000000L;
    if (ts.tv_nsec >= 1000000000L) { ts.tv_sec++; ts.tv_nsec -= 1000000000L; }
    pthread_mutex_lock(&g_mu);
    while (g_qh == g_qt) {
        if (pthread_cond_timedwait(&g_cv, &g_mu, &ts) != 0 && g_qh == g_qt) {
            pthread_mutex_unlock(&g_mu);
            return NVML_ERROR_TIMEOUT;
        }
    }
    ev_t e = g_q[g_qh % 4096];
    g_qh++;
    pthread_mutex_unlock(&g_mu);
    data->device = e.dev >= 0 ? H(e.dev) : (nvmlDevice_t)0;
    data->eventType = e.etype;
    data->eventData = e.data;
    /* gpuInstanceId / computeInstanceId are left as the caller preset them (0xFFFFFFFF) */
    return NVML_SUCCESS;
}
