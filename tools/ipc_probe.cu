// Probe (debug tool): can two PROCESSES on one box share device memory through CUDA IPC and signal each other
// with system-scope stores?  Also times a P2P flag ping-pong and an NCCL send/recv of one halo plane.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/ipc_probe tools/ipc_probe.cu -lnccl && ./tools/ipc_probe
#include <cuda_runtime.h>
#include <nccl.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <unistd.h>
#include <sys/wait.h>
#include <chrono>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("[%d] %s: %s\n", rank, #x, cudaGetErrorString(e)); exit(2); } } while (0)
#define NK(x) do { ncclResult_t e = (x); if (e != ncclSuccess) { printf("[%d] %s: %s\n", rank, #x, ncclGetErrorString(e)); exit(3); } } while (0)

__global__ void k_pingpong(volatile unsigned* mine, unsigned* peer, int rank, int iters) {
	// rank 0 sends odd values, rank 1 answers with even ones
	for (int i = 1; i <= iters; i++) {
		unsigned want = rank == 0 ? 2u * i : 2u * i - 1u;
		if (rank == 0) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(peer), "r"(2u * i - 1u) : "memory"); }
		while (*mine < want) { }
		if (rank == 1) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(peer), "r"(2u * i) : "memory"); }
	}
}
__global__ void k_fill(float* p, size_t n, float v) { size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; if (i < n) p[i] = v; }

int main() {
	int p01[2], p10[2];
	if (pipe(p01) || pipe(p10)) return 1;
	pid_t pid = fork();   // before any CUDA call: a forked child cannot reuse the parent's CUDA context
	int rank = pid == 0 ? 1 : 0;
	int ndev = 0; cudaGetDeviceCount(&ndev);
	if (ndev < 2) { printf("need 2 GPUs, have %d\n", ndev); return 0; }
	int rd = rank == 0 ? p10[0] : p01[0], wr = rank == 0 ? p01[1] : p10[1];
	CK(cudaSetDevice(rank));
	const size_t N = 1 << 20;
	float* buf; unsigned* flag;
	CK(cudaMalloc(&buf, N * 4)); CK(cudaMalloc(&flag, 256)); CK(cudaMemset(flag, 0, 256)); CK(cudaMemset(buf, 0, N * 4));
	cudaIpcMemHandle_t hb, hf, pb, pf;
	CK(cudaIpcGetMemHandle(&hb, buf)); CK(cudaIpcGetMemHandle(&hf, flag));
	ncclUniqueId id; if (rank == 0) NK(ncclGetUniqueId(&id));
	if (rank == 0) { write(wr, &id, sizeof id); } else { read(rd, &id, sizeof id); }
	write(wr, &hb, sizeof hb); write(wr, &hf, sizeof hf);
	read(rd, &pb, sizeof pb); read(rd, &pf, sizeof pf);
	int can = 0; CK(cudaDeviceCanAccessPeer(&can, rank, 1 - rank));
	float* peer_buf; unsigned* peer_flag;
	CK(cudaIpcOpenMemHandle((void**)&peer_buf, pb, cudaIpcMemLazyEnablePeerAccess));
	CK(cudaIpcOpenMemHandle((void**)&peer_flag, pf, cudaIpcMemLazyEnablePeerAccess));
	printf("[%d] canAccessPeer=%d, IPC handles opened\n", rank, can);
	// 1. remote fill + flag
	k_fill<<<(N + 255) / 256, 256>>>(peer_buf, N, (float)(rank + 1));
	CK(cudaDeviceSynchronize());
	char c = 1; write(wr, &c, 1); read(rd, &c, 1);
	float h[4]; CK(cudaMemcpy(h, buf + N - 4, 16, cudaMemcpyDeviceToHost));
	printf("[%d] my buffer was filled by the peer with %g (expect %d)\n", rank, h[3], 2 - rank);
	// 2. flag ping-pong latency
	const int iters = 20000;
	write(wr, &c, 1); read(rd, &c, 1);
	auto t0 = std::chrono::steady_clock::now();
	k_pingpong<<<1, 1>>>(flag, peer_flag, rank, iters);
	CK(cudaDeviceSynchronize());
	double us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count();
	printf("[%d] P2P flag round trip: %.2f us\n", rank, us / iters);
	// 3. NCCL send/recv of one 256^2 float4 plane
	ncclComm_t comm; NK(ncclCommInitRank(&comm, 2, id, rank));
	cudaStream_t st; CK(cudaStreamCreate(&st));
	const size_t plane = 256 * 256 * 4;
	for (int rep = 0; rep < 2; rep++) {
		CK(cudaStreamSynchronize(st));
		auto t1 = std::chrono::steady_clock::now();
		for (int i = 0; i < 200; i++) {
			NK(ncclGroupStart());
			NK(ncclSend(buf, plane, ncclFloat, 1 - rank, comm, st));
			NK(ncclRecv(buf + plane, plane, ncclFloat, 1 - rank, comm, st));
			NK(ncclGroupEnd());
		}
		CK(cudaStreamSynchronize(st));
		double u2 = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t1).count();
		if (rep) printf("[%d] NCCL exchange of a 1 MiB plane: %.1f us each\n", rank, u2 / 200);
	}
	// 4. NCCL allgather 25 MB per rank
	float* big; CK(cudaMalloc(&big, (size_t)2 * 25 * 1000 * 1000));
	for (int rep = 0; rep < 2; rep++) {
		auto t1 = std::chrono::steady_clock::now();
		for (int i = 0; i < 10; i++) NK(ncclAllGather(big + (size_t)rank * 6250000, big, 6250000, ncclFloat, comm, st));
		CK(cudaStreamSynchronize(st));
		double u2 = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t1).count();
		if (rep) printf("[%d] NCCL allgather 2 x 25 MB: %.1f us each\n", rank, u2 / 10);
	}
	ncclCommDestroy(comm);
	if (rank == 0) { int stt; waitpid(pid, &stt, 0); }
	return 0;
}
