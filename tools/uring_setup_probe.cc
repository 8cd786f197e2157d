#include <linux/io_uring.h>
#include <sys/mman.h>
#include <sys/syscall.h>
#include <unistd.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include <mutex>
using clk = std::chrono::steady_clock;
static double ms(clk::time_point a, clk::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); }
int main(int argc, char** argv) {
    int n = argc > 1 ? atoi(argv[1]) : 16; bool serial = argc > 2 && atoi(argv[2]);
    std::vector<double> ph(n * 4);
    std::vector<std::thread> th; std::mutex mu;
    for (int i = 0; i < n; ++i) th.emplace_back([&, i] {
        std::unique_lock<std::mutex> lk(mu, std::defer_lock); if (serial) lk.lock();
        io_uring_params p; memset(&p, 0, sizeof p);
        p.flags = IORING_SETUP_SINGLE_ISSUER | IORING_SETUP_DEFER_TASKRUN | IORING_SETUP_CQSIZE | IORING_SETUP_CLAMP; p.cq_entries = 4096;
        auto t0 = clk::now();
        int fd = syscall(__NR_io_uring_setup, 1024, &p);
        auto t1 = clk::now();
        size_t len = p.cq_off.cqes + p.cq_entries * sizeof(io_uring_cqe);
        void* a = mmap(nullptr, len, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_POPULATE, fd, IORING_OFF_SQ_RING);
        void* b = mmap(nullptr, p.sq_entries * sizeof(io_uring_sqe), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_POPULATE, fd, IORING_OFF_SQES);
        auto t2 = clk::now();
        void* br = mmap(nullptr, 512 * 16, PROT_READ | PROT_WRITE, MAP_ANONYMOUS | MAP_PRIVATE | MAP_POPULATE, -1, 0);
        io_uring_buf_reg reg; memset(&reg, 0, sizeof reg); reg.ring_addr = (uint64_t)br; reg.ring_entries = 512; reg.bgid = 1;
        syscall(__NR_io_uring_register, fd, IORING_REGISTER_PBUF_RING, &reg, 1);
        auto t3 = clk::now();
        syscall(__NR_io_uring_enter, fd, 0, 0, IORING_ENTER_GETEVENTS, nullptr, 0);
        auto t4 = clk::now();
        ph[i*4] = ms(t0,t1); ph[i*4+1] = ms(t1,t2); ph[i*4+2] = ms(t2,t3); ph[i*4+3] = ms(t3,t4);
        (void)a; (void)b;
    });
    for (auto& t : th) t.join();
    double mx[4] = {0,0,0,0}; for (int i = 0; i < n; ++i) for (int k = 0; k < 4; ++k) mx[k] = std::max(mx[k], ph[i*4+k]);
    printf("max setup %.2f  mmap %.2f  register %.2f  first enter %.2f\n", mx[0], mx[1], mx[2], mx[3]);
}
