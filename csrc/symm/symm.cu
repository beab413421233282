// Symmetric-memory manager: one heap per rank, mapped into every peer on the node, plus an NVLS
// multicast alias of it.  This is the transport under the hand-written collectives - the job the
// reference leaves to gloo's TCP pairs (ref:main.py:50) and the NCCL baseline to ncclComm.
//
// Two sharing mechanisms:
//   vmm (default): cuMemCreate + POSIX-FD shareable handle; FDs travel over Unix sockets
//                  (csrc/host/mesh.cpp, SCM_RIGHTS).  Required for cuMulticast* (NVLS).
//   ipc (fallback): cudaMalloc + cudaIpcGetMemHandle, handle bytes travel through the store.
// The driver API is resolved at run time (cudaGetDriverEntryPoint) so the library loads on a
// CPU-only box.
#include <cuda.h>
#include <cudaTypedefs.h>
#include <stdarg.h>
#include <unistd.h>

#include <cstring>
#include <string>
#include <vector>

#include "coll/coll_launch.cuh"
#include "common.cuh"

static thread_local char g_err[1024] = "";
extern "C" const char* dcp_last_error() { return g_err; }
void dcp_set_error(const char* fmt, ...) {
  va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof g_err, fmt, ap); va_end(ap);
}

// DCP_PDL=0 turns programmatic dependent launch off (plain stream order); read once.
bool dcp_pdl_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("DCP_PDL"); v = (e && e[0] == '0') ? 0 : 1; }
  return v != 0;
}

namespace {

struct Drv {
  PFN_cuMemCreate MemCreate = nullptr;
  PFN_cuMemRelease MemRelease = nullptr;
  PFN_cuMemMap MemMap = nullptr;
  PFN_cuMemUnmap MemUnmap = nullptr;
  PFN_cuMemSetAccess MemSetAccess = nullptr;
  PFN_cuMemAddressReserve MemAddressReserve = nullptr;
  PFN_cuMemAddressFree MemAddressFree = nullptr;
  PFN_cuMemExportToShareableHandle MemExport = nullptr;
  PFN_cuMemImportFromShareableHandle MemImport = nullptr;
  PFN_cuMemGetAllocationGranularity MemGran = nullptr;
  PFN_cuMulticastCreate McCreate = nullptr;
  PFN_cuMulticastAddDevice McAddDevice = nullptr;
  PFN_cuMulticastBindMem McBindMem = nullptr;
  PFN_cuMulticastGetGranularity McGran = nullptr;
  PFN_cuGetErrorString GetErrorString = nullptr;
  PFN_cuDeviceGetAttribute DeviceGetAttribute = nullptr;
  bool ok = false;
};

Drv& drv() {
  static Drv d;
  static bool tried = false;
  if (tried) return d;
  tried = true;
  cudaDriverEntryPointQueryResult q;
  bool all = true;
#define LOAD(field, name) \
  if (cudaGetDriverEntryPoint(name, (void**)&d.field, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) { d.field = nullptr; all = false; }
  LOAD(MemCreate, "cuMemCreate") LOAD(MemRelease, "cuMemRelease") LOAD(MemMap, "cuMemMap") LOAD(MemUnmap, "cuMemUnmap")
  LOAD(MemSetAccess, "cuMemSetAccess") LOAD(MemAddressReserve, "cuMemAddressReserve") LOAD(MemAddressFree, "cuMemAddressFree")
  LOAD(MemExport, "cuMemExportToShareableHandle") LOAD(MemImport, "cuMemImportFromShareableHandle")
  LOAD(MemGran, "cuMemGetAllocationGranularity") LOAD(GetErrorString, "cuGetErrorString") LOAD(DeviceGetAttribute, "cuDeviceGetAttribute")
  d.ok = all;
  LOAD(McCreate, "cuMulticastCreate") LOAD(McAddDevice, "cuMulticastAddDevice") LOAD(McBindMem, "cuMulticastBindMem")
  LOAD(McGran, "cuMulticastGetGranularity")
#undef LOAD
  return d;
}

const char* cu_err(CUresult r) {
  const char* s = nullptr;
  if (drv().GetErrorString) drv().GetErrorString(r, &s);
  return s ? s : "unknown CUDA driver error";
}

#define CU_CHECK(expr)                                                                  \
  do {                                                                                  \
    CUresult _r = (expr);                                                               \
    if (_r != CUDA_SUCCESS) {                                                           \
      dcp_set_error("%s failed: %s (%s:%d)", #expr, cu_err(_r), __FILE__, __LINE__);    \
      return -1;                                                                        \
    }                                                                                   \
  } while (0)

struct Symm {
  int rank = 0, world = 1, device = 0;
  bool vmm = true;
  size_t bytes = 0;  // heap size (rounded)
  size_t gran = 2u << 20;
  CUmemGenericAllocationHandle local_handle = 0;
  std::vector<CUmemGenericAllocationHandle> peer_handles;
  std::vector<CUdeviceptr> peer_va;  // peer_va[rank] = own mapping
  std::vector<bool> ipc_opened;
  CUmemGenericAllocationHandle mc_handle = 0;
  CUdeviceptr mc_va = 0;
  bool mc_bound = false;
  int* err_host = nullptr;  // pinned, mapped
  int* err_dev = nullptr;
  DcpDevCtx ctx{};
  bool emu_hint = false;            // set before finalize for members of an emulated world
  struct EmuGroup* emu = nullptr;   // member of an emulated world (all ranks on this GPU): collective launches are recorded, not executed
};

// An emulated world: W heaps on ONE GPU in ONE process, wired to each other as "peers".  Collectives issued on its members are
// queued per rank; dcp_emu_flush() launches slot i of every rank as one cooperative kernel (coll_launch.cuh).
struct EmuGroup {
  int world = 0, device = 0;
  std::vector<Symm*> members;
  std::vector<std::vector<DcpEmuRecord>> pending;   // [rank][slot]
  DcpDevCtx* cs_dev = nullptr;
  unsigned char* args_dev = nullptr;                // kEmuSlots x world x kEmuMaxArgBytes
};
constexpr int kEmuSlots = 64;

int set_access(CUdeviceptr va, size_t bytes, int device) {
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  CU_CHECK(drv().MemSetAccess(va, bytes, &acc, 1));
  return 0;
}

CUmemAllocationProp alloc_prop(int device) {
  CUmemAllocationProp p{};
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = device;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

}  // namespace

extern "C" {

// Whether device `device` supports multicast objects (NVLS).
int dcp_symm_multicast_supported(int device) {
  if (!drv().DeviceGetAttribute) return 0;
  int v = 0;
  if (drv().DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, device) != CUDA_SUCCESS) return 0;
  return v && drv().McCreate != nullptr;
}

// Create this rank's heap.  mode: 0 = vmm (POSIX FD), 1 = ipc (cudaMalloc).  Returns handle or null.
void* dcp_symm_create(int rank, int world, int device, size_t bytes, int mode) {
  if (world > DCP_MAX_RANKS) { dcp_set_error("world size %d exceeds DCP_MAX_RANKS=%d", world, DCP_MAX_RANKS); return nullptr; }
  if (cudaSetDevice(device) != cudaSuccess) { dcp_set_error("cudaSetDevice(%d) failed", device); return nullptr; }
  cudaFree(0);
  auto* s = new Symm;
  s->rank = rank; s->world = world; s->device = device; s->vmm = (mode == 0);
  s->peer_handles.assign(world, 0); s->peer_va.assign(world, 0); s->ipc_opened.assign(world, false);
  if (s->vmm) {
    if (!drv().ok) { dcp_set_error("CUDA driver VMM entry points unavailable"); delete s; return nullptr; }
    CUmemAllocationProp prop = alloc_prop(device);
    size_t gran = 0;
    if (drv().MemGran(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || gran == 0) gran = 2u << 20;
    if (dcp_symm_multicast_supported(device)) {
      CUmulticastObjectProp mp{};
      mp.numDevices = world; mp.size = bytes; mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
      size_t mg = 0;
      if (drv().McGran(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > gran) gran = mg;
    }
    s->bytes = (bytes + gran - 1) / gran * gran;
    s->gran = gran;
    CUresult r = drv().MemCreate(&s->local_handle, s->bytes, &prop, 0);
    if (r != CUDA_SUCCESS) { dcp_set_error("cuMemCreate(%zu) failed: %s", s->bytes, cu_err(r)); delete s; return nullptr; }
    CUdeviceptr va = 0;
    r = drv().MemAddressReserve(&va, s->bytes, gran, 0, 0);
    if (r == CUDA_SUCCESS) r = drv().MemMap(va, s->bytes, 0, s->local_handle, 0);
    if (r != CUDA_SUCCESS || set_access(va, s->bytes, device) != 0) {
      if (r != CUDA_SUCCESS) dcp_set_error("mapping own heap failed: %s", cu_err(r));
      delete s; return nullptr;
    }
    s->peer_va[rank] = va;
  } else {
    s->bytes = (bytes + (2u << 20) - 1) / (2u << 20) * (2u << 20);
    void* p = nullptr;
    if (cudaMalloc(&p, s->bytes) != cudaSuccess) { dcp_set_error("cudaMalloc(%zu) failed", s->bytes); delete s; return nullptr; }
    s->peer_va[rank] = (CUdeviceptr)p;
  }
  cudaMemset((void*)s->peer_va[rank], 0, DcpHeapLayout::kCtlBytes);
  cudaHostAlloc((void**)&s->err_host, sizeof(int), cudaHostAllocMapped);
  *s->err_host = 0;
  cudaHostGetDevicePointer((void**)&s->err_dev, s->err_host, 0);
  cudaDeviceSynchronize();
  return s;
}

size_t dcp_symm_bytes(void* h) { return static_cast<Symm*>(h)->bytes; }
void* dcp_symm_local_ptr(void* h) { auto* s = static_cast<Symm*>(h); return (void*)s->peer_va[s->rank]; }
void* dcp_symm_peer_ptr(void* h, int r) { return (void*)static_cast<Symm*>(h)->peer_va[r]; }
void* dcp_symm_mc_ptr(void* h) { return (void*)static_cast<Symm*>(h)->mc_va; }

// vmm: export own heap as a POSIX FD (caller sends it to peers and closes it).
int dcp_symm_export_fd(void* h) {
  auto* s = static_cast<Symm*>(h);
  int fd = -1;
  CU_CHECK(drv().MemExport(&fd, s->local_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  return fd;
}

// vmm: map peer `r`'s heap from the FD it sent.
int dcp_symm_import_fd(void* h, int r, int fd) {
  auto* s = static_cast<Symm*>(h);
  CU_CHECK(drv().MemImport(&s->peer_handles[r], (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  CUdeviceptr va = 0;
  CU_CHECK(drv().MemAddressReserve(&va, s->bytes, s->gran, 0, 0));
  CU_CHECK(drv().MemMap(va, s->bytes, 0, s->peer_handles[r], 0));
  if (set_access(va, s->bytes, s->device) != 0) return -1;
  s->peer_va[r] = va;
  return 0;
}

// ipc: 64-byte handle of own heap / open a peer's.
int dcp_symm_ipc_handle(void* h, void* out64) {
  auto* s = static_cast<Symm*>(h);
  cudaIpcMemHandle_t hd;
  DCP_CUDA_CHECK(cudaIpcGetMemHandle(&hd, (void*)s->peer_va[s->rank]));
  memcpy(out64, &hd, sizeof hd);
  return 0;
}
int dcp_symm_ipc_open(void* h, int r, const void* in64) {
  auto* s = static_cast<Symm*>(h);
  cudaIpcMemHandle_t hd; memcpy(&hd, in64, sizeof hd);
  void* p = nullptr;
  DCP_CUDA_CHECK(cudaIpcOpenMemHandle(&p, hd, cudaIpcMemLazyEnablePeerAccess));
  s->peer_va[r] = (CUdeviceptr)p; s->ipc_opened[r] = true;
  return 0;
}

// NVLS step 1 (rank 0): create the multicast object, returns an FD to distribute (or -1).
int dcp_symm_mc_create(void* h) {
  auto* s = static_cast<Symm*>(h);
  CUmulticastObjectProp mp{};
  mp.numDevices = s->world; mp.size = s->bytes; mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CU_CHECK(drv().McCreate(&s->mc_handle, &mp));
  int fd = -1;
  CU_CHECK(drv().MemExport(&fd, s->mc_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  return fd;
}
// NVLS step 1' (other ranks): import the multicast object.
int dcp_symm_mc_import(void* h, int fd) {
  auto* s = static_cast<Symm*>(h);
  CU_CHECK(drv().MemImport(&s->mc_handle, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  return 0;
}
// NVLS step 2 (all ranks): add this device.  Every rank must finish this before anyone binds.
int dcp_symm_mc_add_device(void* h) {
  auto* s = static_cast<Symm*>(h);
  CU_CHECK(drv().McAddDevice(s->mc_handle, s->device));
  return 0;
}
// NVLS step 3 (all ranks): bind own heap and map the multicast VA.
int dcp_symm_mc_bind(void* h) {
  auto* s = static_cast<Symm*>(h);
  CU_CHECK(drv().McBindMem(s->mc_handle, 0, s->local_handle, 0, s->bytes, 0));
  s->mc_bound = true;
  CUdeviceptr va = 0;
  CU_CHECK(drv().MemAddressReserve(&va, s->bytes, s->gran, 0, 0));
  CU_CHECK(drv().MemMap(va, s->bytes, 0, s->mc_handle, 0));
  if (set_access(va, s->bytes, s->device) != 0) return -1;
  s->mc_va = va;
  return 0;
}

// Freeze the device context after all peers are mapped.
int dcp_symm_finalize(void* h, double timeout_s, size_t p2p_slot_bytes, size_t staging_bytes, int allow_nvls) {
  auto* s = static_cast<Symm*>(h);
  DcpDevCtx& c = s->ctx;
  c.rank = s->rank; c.world = s->world;
  c.nvls = (allow_nvls && s->mc_va != 0) ? 1 : 0;
  c.mc = c.nvls ? (char*)s->mc_va : nullptr;
  for (int r = 0; r < s->world; ++r) {
    if (!s->peer_va[r]) { dcp_set_error("peer %d not mapped", r); return -1; }
    c.peer[r] = (char*)s->peer_va[r];
  }
  c.timeout_ns = (unsigned long long)(timeout_s * 1e9);
  c.err = s->err_dev;
  c.p2p_blocks = s->emu_hint ? 16 : DCP_P2P_BLOCKS;
  c.p2p_slot_off = DcpHeapLayout::kCtlBytes;
  c.p2p_slot_bytes = p2p_slot_bytes;
  c.staging_off = c.p2p_slot_off + p2p_slot_bytes * DCP_MAX_RANKS;
  c.staging_bytes = staging_bytes;
  if (c.staging_off + staging_bytes > s->bytes) { dcp_set_error("heap too small for staging"); return -1; }
  return 0;
}

size_t dcp_symm_user_offset(void* h) { auto* s = static_cast<Symm*>(h); return s->ctx.staging_off + s->ctx.staging_bytes; }
size_t dcp_symm_staging_offset(void* h) { return static_cast<Symm*>(h)->ctx.staging_off; }
size_t dcp_symm_staging_bytes(void* h) { return static_cast<Symm*>(h)->ctx.staging_bytes; }
int dcp_symm_has_nvls(void* h) { return static_cast<Symm*>(h)->ctx.nvls; }
const DcpDevCtx* dcp_symm_ctx(void* h) { return &static_cast<Symm*>(h)->ctx; }

// Non-zero once any kernel's bounded wait expired (a peer died or the ranks diverged).
int dcp_symm_error_flag(void* h) { return *((volatile int*)static_cast<Symm*>(h)->err_host); }
void dcp_symm_clear_error(void* h) { *((volatile int*)static_cast<Symm*>(h)->err_host) = 0; }

void dcp_symm_destroy(void* h) {
  auto* s = static_cast<Symm*>(h);
  cudaSetDevice(s->device);
  cudaDeviceSynchronize();
  if (s->vmm) {
    if (s->mc_va) { drv().MemUnmap(s->mc_va, s->bytes); drv().MemAddressFree(s->mc_va, s->bytes); }
    if (s->mc_handle) drv().MemRelease(s->mc_handle);
    for (int r = 0; r < s->world; ++r) {
      if (!s->peer_va[r]) continue;
      drv().MemUnmap(s->peer_va[r], s->bytes);
      drv().MemAddressFree(s->peer_va[r], s->bytes);
      if (r != s->rank && s->peer_handles[r]) drv().MemRelease(s->peer_handles[r]);
    }
    if (s->local_handle) drv().MemRelease(s->local_handle);
  } else {
    for (int r = 0; r < s->world; ++r)
      if (r != s->rank && s->ipc_opened[r]) cudaIpcCloseMemHandle((void*)s->peer_va[r]);
    cudaFree((void*)s->peer_va[s->rank]);
  }
  if (s->err_host) cudaFreeHost(s->err_host);
  delete s;
}

}  // extern "C"

// ---- emulated world (single-GPU test tier) -----------------------------------------------------------
bool dcp_emu_record(void* h, const DcpEmuRecord& r) {
  auto* s = static_cast<Symm*>(h);
  if (s->emu == nullptr) return false;
  s->emu->pending[s->rank].push_back(r);
  return true;
}

extern "C" {

// Create `world` heaps of heap_bytes each on `device`, every one seeing the others as its peers (plain device pointers: same
// process, same GPU), no multicast.  Returns the group; dcp_emu_rank(group, r) is an ordinary heap handle for every other
// entry point of this library - except that collectives issued on it are deferred until dcp_emu_flush().
void* dcp_emu_create(int world, int device, size_t heap_bytes, double timeout_s, size_t p2p_slot_bytes, size_t staging_bytes) {
  if (world < 1 || world > DCP_MAX_RANKS) { dcp_set_error("emulated world size %d not in [1, %d]", world, DCP_MAX_RANKS); return nullptr; }
  auto* g = new EmuGroup;
  g->world = world; g->device = device;
  g->pending.resize(world);
  for (int r = 0; r < world; ++r) {
    auto* s = static_cast<Symm*>(dcp_symm_create(r, world, device, heap_bytes, 1));
    if (!s) { for (auto* m : g->members) dcp_symm_destroy(m); delete g; return nullptr; }
    g->members.push_back(s);
  }
  std::vector<DcpDevCtx> cs(world);
  for (int r = 0; r < world; ++r) {
    Symm* s = g->members[r];
    for (int p = 0; p < world; ++p) s->peer_va[p] = g->members[p]->peer_va[p];
    s->emu_hint = true;
    if (dcp_symm_finalize(s, timeout_s, p2p_slot_bytes, staging_bytes, 0) != 0) { for (auto* m : g->members) dcp_symm_destroy(m); delete g; return nullptr; }
    s->emu = g;
    cs[r] = s->ctx;
  }
  if (cudaMalloc((void**)&g->cs_dev, sizeof(DcpDevCtx) * world) != cudaSuccess ||
      cudaMalloc((void**)&g->args_dev, (size_t)kEmuSlots * world * kEmuMaxArgBytes) != cudaSuccess ||
      cudaMemcpy(g->cs_dev, cs.data(), sizeof(DcpDevCtx) * world, cudaMemcpyHostToDevice) != cudaSuccess) {
    dcp_set_error("emulated world: device allocation failed"); return nullptr;
  }
  return g;
}

void* dcp_emu_rank(void* group, int r) {
  auto* g = static_cast<EmuGroup*>(group);
  return (r >= 0 && r < g->world) ? g->members[r] : nullptr;
}

int dcp_emu_pending(void* group) {
  auto* g = static_cast<EmuGroup*>(group);
  size_t n = 0;
  for (auto& q : g->pending) n = q.size() > n ? q.size() : n;
  return (int)n;
}

// Launch everything the ranks have queued: slot i of every rank becomes ONE cooperative kernel of (grid, world) CTAs on
// `stream`.  The ranks must have queued the same collectives in the same order (same kernel, same grid) - as they would have
// to on N GPUs; point-to-point slots may be empty on ranks that take no part.  Returns -1 (and drops the queue) on a mismatch.
int dcp_emu_flush(void* group, cudaStream_t stream) {
  auto* g = static_cast<EmuGroup*>(group);
  const int W = g->world;
  const int n = dcp_emu_pending(group);
  int rc = 0;
  std::vector<unsigned char> host((size_t)W * kEmuMaxArgBytes);
  for (int i = 0; i < n && rc == 0; ++i) {
    const DcpEmuRecord* ref = nullptr;
    for (int r = 0; r < W; ++r) if ((int)g->pending[r].size() > i) { ref = &g->pending[r][i]; break; }
    std::fill(host.begin(), host.end(), 0);
    for (int r = 0; r < W && rc == 0; ++r) {
      if ((int)g->pending[r].size() <= i) {
        if (!ref->idle_ok) { dcp_set_error("emulated world: rank %d issued %zu collectives, another rank at least %d", r, g->pending[r].size(), i + 1); rc = -1; }
        continue;
      }
      const DcpEmuRecord& q = g->pending[r][i];
      if (q.launch != ref->launch || q.grid != ref->grid || q.arg_bytes != ref->arg_bytes) {
        dcp_set_error("emulated world: the ranks issued different collectives in slot %d (kernel or grid differ)", i); rc = -1; continue;
      }
      memcpy(host.data() + (size_t)r * q.arg_bytes, q.args, q.arg_bytes);
    }
    if (rc != 0) break;
    if (i > 0 && i % kEmuSlots == 0) cudaStreamSynchronize(stream);       // the argument ring is about to wrap
    unsigned char* dst = g->args_dev + (size_t)(i % kEmuSlots) * W * kEmuMaxArgBytes;
    if (cudaMemcpyAsync(dst, host.data(), (size_t)W * ref->arg_bytes, cudaMemcpyHostToDevice, stream) != cudaSuccess) { dcp_set_error("emulated world: argument upload failed"); rc = -1; break; }
    cudaError_t e = ref->launch(g->cs_dev, dst, ref->grid, W, stream);
    if (e != cudaSuccess) { dcp_set_error("emulated world: cooperative launch failed: %s", cudaGetErrorString(e)); rc = -1; }
  }
  for (auto& q : g->pending) q.clear();
  return rc;
}

void dcp_emu_destroy(void* group) {
  auto* g = static_cast<EmuGroup*>(group);
  cudaSetDevice(g->device);
  cudaDeviceSynchronize();
  for (auto* m : g->members) {
    for (int p = 0; p < g->world; ++p) if (p != m->rank) m->peer_va[p] = 0;     // peers belong to their own handles
    dcp_symm_destroy(m);
  }
  if (g->cs_dev) cudaFree(g->cs_dev);
  if (g->args_dev) cudaFree(g->args_dev);
  delete g;
}

}  // extern "C"
