// TEST INFRASTRUCTURE ONLY (oracle/): Linux implementation of the slice of the reference's platform
// layer "PL" (declared in /root/reference/Source/PL/pl_utils.h:6-221, implemented upstream only for
// Win32 in PL/Win32/pl_utils_win32.cpp) that the hot-path sources need in order to link.
// This file is #included by ref_harness.cpp (unity TU). Nothing in the product links against it.
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <stdarg.h>
#include <pthread.h>
#include <unistd.h>
#include <time.h>
#include <sys/syscall.h>

// pl_utils_win32.cpp:84-93 uses HeapAlloc(HEAP_ZERO_MEMORY): allocations come back zeroed and
// HeapReAlloc zero-fills the grown tail. The builder relies on that (kd_tree.cpp:22, FDBuffer::allocate).
// A 16-byte size header keeps the zero-extension exact without relying on malloc slack.
void* pl_buffer_alloc(size_t size)
{
	char* p = (char*)calloc(1, size + 16);
	*(size_t*)p = size;
	return p + 16;
}
void* pl_buffer_resize(void* block, size_t new_size)
{
	if (!block) return pl_buffer_alloc(new_size);
	char* base = (char*)block - 16;
	size_t old = *(size_t*)base;
	char* p = (char*)realloc(base, new_size + 16);
	*(size_t*)p = new_size;
	if (new_size > old) memset(p + 16 + old, 0, new_size - old);
	return p + 16;
}
void pl_buffer_free(void* buffer) { if (buffer) free((char*)buffer - 16); }
void pl_buffer_copy(void* destination, void* from, size_t length) { if (length) memcpy(destination, from, length); }
void pl_buffer_move(void* destination, void* source, size_t length) { if (length) memmove(destination, source, length); }
void pl_buffer_set(void* buffer, int32 value_to_set, size_t n) { memset(buffer, value_to_set, n); }

struct ShimThread
{
	pthread_t tid;
	ThreadProc proc;
	void* data;
	volatile int done;
	int joined;
};
static void* shim_thread_main(void* p)
{
	ShimThread* t = (ShimThread*)p;
	t->proc(t->data);
	__sync_synchronize();
	t->done = 1;
	return 0;
}
ThreadHandle pl_create_thread(ThreadProc proc, void* data)
{
	ShimThread* t = (ShimThread*)calloc(1, sizeof(ShimThread));
	t->proc = proc;
	t->data = data;
	pthread_create(&t->tid, 0, shim_thread_main, t);
	ThreadHandle h;
	h.thread_handle = t;
	return h;
}
void pl_close_thread(const ThreadHandle* handle) { (void)handle; }
static double shim_now_ms()
{
	struct timespec ts;
	clock_gettime(CLOCK_MONOTONIC, &ts);
	return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}
// Win32 WaitForMultipleObjects(all): 0 (WAIT_OBJECT_0) when every thread finished, non-zero on timeout
// (pl_utils_win32.cpp:28-31).
b32 pl_wait_for_all_threads(uint32 n, const ThreadHandle* handles, uint32 timeout_in_ms)
{
	double deadline = shim_now_ms() + (double)timeout_in_ms;
	for (;;)
	{
		uint32 done = 0;
		for (uint32 i = 0; i < n; i++) done += ((ShimThread*)handles[i].thread_handle)->done ? 1 : 0;
		if (done == n) break;
		if (timeout_in_ms != 0xffffffffu && shim_now_ms() >= deadline) return 258; // WAIT_TIMEOUT
		usleep(100);
	}
	for (uint32 i = 0; i < n; i++)
	{
		ShimThread* t = (ShimThread*)handles[i].thread_handle;
		if (!t->joined) { pthread_join(t->tid, 0); t->joined = 1; }
	}
	return 0;
}
b32 pl_wait_for_thread(const ThreadHandle* handle, uint32 timeout_in_ms) { return pl_wait_for_all_threads(1, handle, timeout_in_ms); }
void pl_sleep_thread(uint32 ms) { usleep(ms * 1000u); }
uint32 pl_get_thread_id() { return (uint32)syscall(SYS_gettid); }
// Upstream: __rdtsc() * thread id (pl_utils_win32.cpp:10-15) — non-deterministic; irrelevant at bounce_limit 1.
uint64 pl_get_hardware_entropy() { return 0x853c49e6748fea9bULL * (uint64)pl_get_thread_id(); }
uint64 pl_get_tsc() { return __rdtsc(); }

b32 pl_get_file_handle(char* path, void** handle) { FILE* f = fopen(path, "rb"); *handle = f; return f != 0; }
uint64 pl_get_file_size(void* handle)
{
	FILE* f = (FILE*)handle;
	long cur = ftell(f);
	fseek(f, 0, SEEK_END);
	long sz = ftell(f);
	fseek(f, cur, SEEK_SET);
	return (uint64)sz;
}
b32 pl_load_file_into(void* handle, void* block, uint32 file_size) { return fread(block, 1, file_size, (FILE*)handle) == file_size; }
// CreateFileA(..., CREATE_NEW, ...) (PL/Win32/pl_utils_win32.cpp:143-158): fails when the file already exists
b32 pl_create_file(void** file_handle, char* path) { FILE* f = fopen(path, "wbx"); *file_handle = f; return f != 0; }
b32 pl_append_to_file(void* file_handle, void* block, int32 n) { return fwrite(block, 1, n, (FILE*)file_handle) == (size_t)n; }
b32 pl_close_file_handle(void* file_handle) { return fclose((FILE*)file_handle) == 0; }
void pl_throw_error_box(const char* error) { fprintf(stderr, "PL error: %s\n", error); }
void pl_debug_print(const char* format, ...) { (void)format; }
void pl_format_print(char* buffer, uint32 buffer_size, const char* format, ...)
{
	va_list ap;
	va_start(ap, format);
	vsnprintf(buffer, buffer_size, format, ap);
	va_end(ap);
}
