// fs3_divergence_no_alloc: PrintDivergence (S3D.cpp:1095-1138) on caller-supplied device grids WITHOUT any allocation.
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is a variant of the divergence metric
// of fs3_get_divergence (fs3_hooks.cu), whose block partials live in the handle's cudaMalloc'ed staging buffer; here they
// live in a file-scope __device__ array of fixed size, which every GPU that loads the library holds until the process exits.
//
// The array's NAME is a host-side stand-in in host code, not a device address: its device address is taken once per device,
// with that device current, through cudaGetSymbolAddress, and that pointer is what the kernels get.
// All calls share the array and it keeps the last call's contents.  Nothing in it has to start at zero: a call's first kernel
// overwrites every partial its second kernel reads (exactly `blocks` pairs), so stale contents are never consumed.  Calls
// that overlap on the GPU (different streams, or different host threads) RACE on it: serialise them on one stream.
#include "fs3_internal.cuh"
#include <mutex>

void fs3i_set_global_error(const std::string& msg);   // fs3_api.cu: error text for entry points without a handle

#define NA_THREADS 256
#define NA_MAX_BLOCKS (FS3_NO_ALLOC_MAX_CELLS / NA_THREADS)
// [2 * b] = sum of squares of block b, [2 * b + 1] = its maximum: 65 536 pairs = 1 MiB
static __device__ __align__(256) double g_na_scratch[2 * NA_MAX_BLOCKS];

__global__ void __launch_bounds__(NA_THREADS) k_na_divergence(const float* __restrict__ U, const float* __restrict__ V, const float* __restrict__ W,
                                                              const float* __restrict__ phi, int nx, int ny, int nz, double* __restrict__ part) {
	const size_t c = (size_t)blockIdx.x * NA_THREADS + threadIdx.x, nC = (size_t)nx * ny * nz;
	double sq = 0.0, mx = 0.0;
	if (c < nC && phi[c] < 0.0f) {
		const int x = (int)(c % nx), y = (int)((c / nx) % ny), z = (int)(c / ((size_t)nx * ny));
		const size_t iu = (size_t)x + (size_t)(nx + 1) * ((size_t)y + (size_t)ny * z);
		const size_t iv = (size_t)x + (size_t)nx * ((size_t)y + (size_t)(ny + 1) * z);
		// velUp - velDown + velRight - velLeft + velFwds - velBkwds, left to right in fp32 (S3D.cpp:1118)
		float dv = __fsub_rn(V[iv + nx], V[iv]);
		dv = __fadd_rn(dv, U[iu + 1]); dv = __fsub_rn(dv, U[iu]);
		dv = __fadd_rn(dv, W[c + (size_t)nx * ny]); dv = __fsub_rn(dv, W[c]);
		mx = dv > 0.0f ? (double)dv : 0.0;
		sq = (double)__fmul_rn(dv, dv);
	}
	__shared__ double s_sq[NA_THREADS], s_mx[NA_THREADS];
	s_sq[threadIdx.x] = sq; s_mx[threadIdx.x] = mx;
	__syncthreads();
	for (int o = NA_THREADS / 2; o > 0; o >>= 1) {
		if (threadIdx.x < o) { s_sq[threadIdx.x] += s_sq[threadIdx.x + o]; s_mx[threadIdx.x] = fmax(s_mx[threadIdx.x], s_mx[threadIdx.x + o]); }
		__syncthreads();
	}
	if (threadIdx.x == 0) { part[2 * (size_t)blockIdx.x] = s_sq[0]; part[2 * (size_t)blockIdx.x + 1] = s_mx[0]; }
}
// out[0] = max divergence, out[1] = sqrt(sum of squares); partials added in a fixed order (the same order as fs3_get_divergence)
__global__ void __launch_bounds__(NA_THREADS) k_na_final(const double* __restrict__ part, unsigned nblocks, double* __restrict__ out) {
	__shared__ double s_sq[NA_THREADS], s_mx[NA_THREADS];
	double sq = 0.0, mx = 0.0;
	for (unsigned b = threadIdx.x; b < nblocks; b += NA_THREADS) { sq += part[2 * (size_t)b]; mx = fmax(mx, part[2 * (size_t)b + 1]); }
	s_sq[threadIdx.x] = sq; s_mx[threadIdx.x] = mx;
	__syncthreads();
	for (int o = NA_THREADS / 2; o > 0; o >>= 1) {
		if (threadIdx.x < o) { s_sq[threadIdx.x] += s_sq[threadIdx.x + o]; s_mx[threadIdx.x] = fmax(s_mx[threadIdx.x], s_mx[threadIdx.x + o]); }
		__syncthreads();
	}
	if (threadIdx.x == 0) { out[0] = s_mx[0]; out[1] = sqrt(s_sq[0]); }
}

// device address of the scratch array, taken once per device with that device current
#define NA_MAX_DEVICES 64
static double* g_na_ptr[NA_MAX_DEVICES];
static std::mutex g_na_mutex;

static int na_fail(int code, const std::string& msg) { fs3i_set_global_error(msg); return code; }

extern "C" int fs3_divergence_no_alloc(int device, void* stream, int nx, int ny, int nz, const float* dev_U, const float* dev_V, const float* dev_W,
                                       const float* dev_phi, double* dev_out) {
	if (!dev_U || !dev_V || !dev_W || !dev_phi || !dev_out) return na_fail(FS3_ERR_INVALID, "fs3_divergence_no_alloc: null pointer argument");
	if (nx < 1 || ny < 1 || nz < 1) return na_fail(FS3_ERR_INVALID, "fs3_divergence_no_alloc: grid extents must be positive");
	const unsigned long long cells = (unsigned long long)nx * (unsigned long long)ny * (unsigned long long)nz;
	if (cells > (unsigned long long)FS3_NO_ALLOC_MAX_CELLS)
		return na_fail(FS3_ERR_INVALID, "fs3_divergence_no_alloc: grid has " + std::to_string(cells) + " cells, the fixed scratch array holds the partial sums of at most " +
		                                    std::to_string((unsigned long long)FS3_NO_ALLOC_MAX_CELLS) + " (FS3_NO_ALLOC_MAX_CELLS = 256^3); use fs3_get_divergence");
	if (device < 0 || device >= NA_MAX_DEVICES) return na_fail(FS3_ERR_INVALID, "fs3_divergence_no_alloc: device ordinal out of range");
	cudaError_t e = cudaSetDevice(device);
	if (e != cudaSuccess) return na_fail(FS3_ERR_CUDA, std::string("fs3_divergence_no_alloc: cudaSetDevice: ") + cudaGetErrorString(e));
	double* scratch;
	{
		std::lock_guard<std::mutex> lock(g_na_mutex);
		if (!g_na_ptr[device]) {
			// set-up, once per device: the array's name is only a host-side stand-in -- ask the runtime for its address on THIS device
			void* p = nullptr;
			e = cudaGetSymbolAddress(&p, g_na_scratch);
			if (e != cudaSuccess || !p) return na_fail(FS3_ERR_CUDA, std::string("fs3_divergence_no_alloc: cudaGetSymbolAddress (is this an sm_100 device?): ") + cudaGetErrorString(e));
			g_na_ptr[device] = (double*)p;
		}
		scratch = g_na_ptr[device];
	}
	const unsigned nb = (unsigned)((cells + NA_THREADS - 1) / NA_THREADS);
	cudaStream_t st = (cudaStream_t)stream;
	k_na_divergence<<<nb, NA_THREADS, 0, st>>>(dev_U, dev_V, dev_W, dev_phi, nx, ny, nz, scratch);
	k_na_final<<<1, NA_THREADS, 0, st>>>(scratch, nb, dev_out);
	e = cudaGetLastError();
	if (e != cudaSuccess) return na_fail(FS3_ERR_CUDA, std::string("fs3_divergence_no_alloc: kernel launch: ") + cudaGetErrorString(e));
	return FS3_OK;
}
