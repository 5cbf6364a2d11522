// Native parameter-server runtime (libpsb200_ps.so) — the CPU/TCP transport.
//
// One instance runs inside every task process (like tf.train.Server, encoder:96-98); in a
// `--job_name=ps` process it is the whole job: it stores the variables placed on this ps task,
// applies optimizer updates where the variable lives (TF colocates Apply* with its variable),
// hosts the shared-name FIFO queues (done-queues encoder:54-63, sync token queue) and the
// per-variable gradient accumulators of sync mode (SyncReplicasOptimizer).
//
// Concurrency model: one thread per connection.  Async applies are deliberately lock-free
// (TF's use_locking=False: Hogwild) — two workers may interleave on one variable.
// Queues and accumulators block on condition variables, never on a global lock.
//
// Wire format (little endian):
//   request : u32 magic 'PSB1' | u32 op | u32 name_len | u64 payload_len | name | payload
//   response: i32 status (0 ok, <0 error) | u64 payload_len | payload
#include <arpa/inet.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

enum Op : uint32_t {
    OP_PING = 1, OP_VAR_DECLARE = 2, OP_VAR_ASSIGN = 3, OP_VAR_READ = 4, OP_VAR_UNINIT = 5,
    OP_APPLY = 6, OP_VAR_ADD_I32 = 7, OP_VAR_SCALE_F32 = 8, OP_Q_CREATE = 9, OP_Q_ENQUEUE = 10,
    OP_Q_DEQUEUE = 11, OP_Q_SIZE = 12, OP_ACC_DECLARE = 13, OP_ACC_APPLY = 14,
    OP_ACC_TAKE_APPLY = 15, OP_ACC_SET_STEP = 16, OP_SHUTDOWN = 17, OP_VAR_LIST = 18,
    OP_STATS = 19,
};

enum Status : int32_t {
    ST_OK = 0, ST_BAD_REQUEST = -1, ST_UNINITIALIZED = -2, ST_TIMEOUT = -3, ST_NOT_FOUND = -4,
    ST_SHAPE = -5, ST_STOPPED = -6,
};

constexpr uint32_t kMagic = 0x31425350;  // "PSB1"
constexpr uint64_t kMaxNumel = 1ull << 32;  // elements per variable a declare may ask for (16 GiB of fp32)

struct Variable {
    uint32_t dtype = 0;  // 0 f32, 1 i32
    std::vector<uint64_t> dims;
    std::vector<uint8_t> data;
    std::atomic<bool> initialized{false};
    size_t numel() const { return data.size() / 4; }
};

struct Queue {
    std::mutex mu;
    std::condition_variable not_empty, not_full;
    std::deque<int32_t> items;
    size_t capacity = 1;
};

struct Accumulator {
    std::mutex mu;
    std::condition_variable cv;
    std::vector<float> sum;
    uint32_t count = 0;
    int64_t global_step = 0;
    uint64_t dropped = 0;
};

#pragma pack(push, 1)
struct ApplyHdr {
    uint32_t kind;  // 0 SGD, 1 RMSProp, 2 Adam
    float lr, h1, h2, eps;
    float b1p, b2p;  // Adam: current beta1^t, beta2^t (read by the worker from their owner)
    uint32_t nslots;
};
#pragma pack(pop)

struct Server {
    int listen_fd = -1;
    int port = 0;
    std::atomic<bool> stopping{false};
    std::thread accept_thread;
    std::shared_mutex reg_mu;
    std::map<std::string, std::unique_ptr<Variable>> vars;
    std::map<std::string, std::unique_ptr<Queue>> queues;
    std::map<std::string, std::unique_ptr<Accumulator>> accs;
    std::atomic<uint64_t> n_requests{0}, n_applies{0}, bytes_in{0}, bytes_out{0};
    std::mutex conn_mu;
    std::vector<int> conn_fds;

    Variable* find_var(const std::string& n) {
        std::shared_lock<std::shared_mutex> l(reg_mu);
        auto it = vars.find(n);
        return it == vars.end() ? nullptr : it->second.get();
    }
    Queue* find_queue(const std::string& n) {
        std::shared_lock<std::shared_mutex> l(reg_mu);
        auto it = queues.find(n);
        return it == queues.end() ? nullptr : it->second.get();
    }
    Accumulator* find_acc(const std::string& n) {
        std::shared_lock<std::shared_mutex> l(reg_mu);
        auto it = accs.find(n);
        return it == accs.end() ? nullptr : it->second.get();
    }
};

bool read_full(int fd, void* buf, size_t n) {
    uint8_t* p = static_cast<uint8_t*>(buf);
    while (n > 0) {
        ssize_t r = ::recv(fd, p, n, 0);
        if (r <= 0) return false;
        p += r;
        n -= static_cast<size_t>(r);
    }
    return true;
}

bool write_full(int fd, const void* buf, size_t n) {
    const uint8_t* p = static_cast<const uint8_t*>(buf);
    while (n > 0) {
        ssize_t r = ::send(fd, p, n, MSG_NOSIGNAL);
        if (r <= 0) return false;
        p += r;
        n -= static_cast<size_t>(r);
    }
    return true;
}

bool respond(Server* s, int fd, int32_t status, const void* payload = nullptr, uint64_t len = 0) {
    uint8_t hdr[12];
    memcpy(hdr, &status, 4);
    memcpy(hdr + 4, &len, 8);
    if (!write_full(fd, hdr, 12)) return false;
    if (len && !write_full(fd, payload, len)) return false;
    s->bytes_out += 12 + len;
    return true;
}

// The optimizer arithmetic of TF 1.11's dense Apply* kernels.
void apply_dense(const ApplyHdr& h, float* w, const float* g, float* s0, float* s1, size_t n, float scale) {
    if (h.kind == 0) {
        for (size_t i = 0; i < n; ++i) w[i] -= h.lr * (g[i] * scale);
    } else if (h.kind == 1) {
        for (size_t i = 0; i < n; ++i) {
            const float gi = g[i] * scale;
            float ms = h.h1 * s0[i] + (1.0f - h.h1) * gi * gi;
            float mom = h.h2 * s1[i] + h.lr * gi / std::sqrt(ms + h.eps);
            s0[i] = ms;
            s1[i] = mom;
            w[i] -= mom;
        }
    } else {
        const float lr_t = h.lr * std::sqrt(1.0f - h.b2p) / (1.0f - h.b1p);
        for (size_t i = 0; i < n; ++i) {
            const float gi = g[i] * scale;
            float m = s0[i] + (1.0f - h.h1) * (gi - s0[i]);
            float v = s1[i] + (1.0f - h.h2) * (gi * gi - s1[i]);
            s0[i] = m;
            s1[i] = v;
            w[i] -= lr_t * m / (std::sqrt(v) + h.eps);
        }
    }
}

// Parses `ApplyHdr | nslots x (u32 len, bytes)`; returns bytes consumed or 0 on error.
size_t parse_apply(Server* s, const std::vector<uint8_t>& p, ApplyHdr* h, Variable** slots) {
    if (p.size() < sizeof(ApplyHdr)) return 0;
    memcpy(h, p.data(), sizeof(ApplyHdr));
    size_t off = sizeof(ApplyHdr);
    // SGD carries no slots, RMSProp and Adam exactly two (rms, momentum | m, v)
    if (h->kind > 2 || h->nslots != (h->kind == 0 ? 0u : 2u)) return 0;
    for (uint32_t k = 0; k < h->nslots; ++k) {
        if (off + 4 > p.size()) return 0;
        uint32_t len;
        memcpy(&len, p.data() + off, 4);
        off += 4;
        if (off + len > p.size()) return 0;
        std::string name(reinterpret_cast<const char*>(p.data() + off), len);
        off += len;
        slots[k] = s->find_var(name);
        if (!slots[k] || slots[k]->dtype != 0) return 0;
    }
    return off;
}

// Every slot a kind needs is present and as large as the variable it belongs to.
bool slots_fit(const ApplyHdr& h, Variable* const* slots, const Variable* v) {
    if (v->dtype != 0) return false;
    for (uint32_t k = 0; k < h.nslots; ++k)
        if (!slots[k] || slots[k]->data.size() != v->data.size()) return false;
    return true;
}

void handle(Server* s, int fd) {
    int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
    std::vector<uint8_t> payload;
    while (!s->stopping.load()) {
        uint8_t hdr[20];
        if (!read_full(fd, hdr, 20)) break;
        uint32_t magic, op, name_len;
        uint64_t plen;
        memcpy(&magic, hdr, 4);
        memcpy(&op, hdr + 4, 4);
        memcpy(&name_len, hdr + 8, 4);
        memcpy(&plen, hdr + 12, 8);
        if (magic != kMagic || name_len > 4096 || plen > (1ull << 33)) break;
        std::string name(name_len, '\0');
        if (name_len && !read_full(fd, &name[0], name_len)) break;
        payload.resize(plen);
        if (plen && !read_full(fd, payload.data(), plen)) break;
        s->n_requests++;
        s->bytes_in += 20 + name_len + plen;
        bool ok = true;
        try {
        switch (op) {
            case OP_PING:
                ok = respond(s, fd, ST_OK);
                break;
            case OP_VAR_DECLARE: {
                if (plen < 8) { ok = respond(s, fd, ST_BAD_REQUEST); break; }
                uint32_t dtype, ndim;
                memcpy(&dtype, payload.data(), 4);
                memcpy(&ndim, payload.data() + 4, 4);
                if (plen != 8 + 8ull * ndim || ndim > 8) { ok = respond(s, fd, ST_BAD_REQUEST); break; }
                std::vector<uint64_t> dims(ndim);
                uint64_t numel = 1;
                bool too_big = false;
                for (uint32_t i = 0; i < ndim; ++i) {
                    memcpy(&dims[i], payload.data() + 8 + 8 * i, 8);
                    if (dims[i] > kMaxNumel || (dims[i] != 0 && numel > kMaxNumel / dims[i])) too_big = true;
                    else numel *= dims[i];
                }
                if (too_big || dtype > 1) { ok = respond(s, fd, ST_BAD_REQUEST); break; }
                uint32_t inited = 0;
                int32_t st = ST_OK;
                {
                    std::unique_lock<std::shared_mutex> l(s->reg_mu);
                    auto it = s->vars.find(name);
                    if (it == s->vars.end()) {
                        auto v = std::make_unique<Variable>();
                        v->dtype = dtype;
                        v->dims = dims;
                        v->data.assign(numel * 4, 0);
                        s->vars.emplace(name, std::move(v));
                    } else {
                        if (it->second->dims != dims || it->second->dtype != dtype) st = ST_SHAPE;
                        inited = it->second->initialized.load() ? 1 : 0;
                    }
                }
                ok = respond(s, fd, st, &inited, 4);
                break;
            }
            case OP_VAR_ASSIGN: {
                Variable* v = s->find_var(name);
                if (!v) { ok = respond(s, fd, ST_NOT_FOUND); break; }
                if (plen != v->data.size()) { ok = respond(s, fd, ST_SHAPE); break; }
                memcpy(v->data.data(), payload.data(), plen);
                v->initialized.store(true);
                ok = respond(s, fd, ST_OK);
                break;
            }
            case OP_VAR_READ: {
                Variable* v = s->find_var(name);
                if (!v) { ok = respond(s, fd, ST_NOT_FOUND); break; }
                if (!v->initialized.load()) { ok = respond(s, fd, ST_UNINITIALIZED); break; }
                ok = respond(s, fd, ST_OK, v->data.data(), v->data.size());
                break;
            }
            case OP_VAR_UNINIT: {
                uint32_t n = 0;
                {
                    std::shared_lock<std::shared_mutex> l(s->reg_mu);
                    for (auto& kv : s->vars)
                        if (!kv.second->initialized.load()) ++n;
                }
                ok = respond(s, fd, ST_OK, &n, 4);
                break;
            }
            case OP_VAR_LIST: {
                std::string out;
                {
                    std::shared_lock<std::shared_mutex> l(s->reg_mu);
                    for (auto& kv : s->vars) {
                        out += kv.first;
                        out += '\n';
                    }
                }
                ok = respond(s, fd, ST_OK, out.data(), out.size());
                break;
            }
            case OP_APPLY: {
                Variable* v = s->find_var(name);
                if (!v) { ok = respond(s, fd, ST_NOT_FOUND); break; }
                if (!v->initialized.load()) { ok = respond(s, fd, ST_UNINITIALIZED); break; }
                ApplyHdr h;
                Variable* slots[2] = {nullptr, nullptr};
                size_t off = parse_apply(s, payload, &h, slots);
                if (!off || payload.size() - off != v->data.size()) { ok = respond(s, fd, ST_BAD_REQUEST); break; }
                if (!slots_fit(h, slots, v)) { ok = respond(s, fd, ST_SHAPE); break; }
                // gradient bytes may be unaligned inside the payload (slot names have any length):
                // use them in place when they are float-aligned, copy out otherwise
                const uint8_t* gbytes = payload.data() + off;
                std::vector<float> gcopy;
                if (reinterpret_cast<uintptr_t>(gbytes) % alignof(float) != 0) {
                    gcopy.resize(v->numel());
                    memcpy(gcopy.data(), gbytes, v->data.size());
                    gbytes = reinterpret_cast<const uint8_t*>(gcopy.data());
                }
                apply_dense(h, reinterpret_cast<float*>(v->data.data()), reinterpret_cast<const float*>(gbytes),
                            slots[0] ? reinterpret_cast<float*>(slots[0]->data.data()) : nullptr,
                            slots[1] ? reinterpret_cast<float*>(slots[1]->data.data()) : nullptr, v->numel(), 1.0f);
                s->n_applies++;
                ok = respond(s, fd, ST_OK);
                break;
            }
            case OP_VAR_ADD_I32: {
                Variable* v = s->find_var(name);
                if (!v || v->dtype != 1 || plen != 4) { ok = respond(s, fd, v ? ST_BAD_REQUEST : ST_NOT_FOUND); break; }
                int32_t delta;
                memcpy(&delta, payload.data(), 4);
                int32_t nv = __atomic_add_fetch(reinterpret_cast<int32_t*>(v->data.data()), delta, __ATOMIC_SEQ_CST);
                ok = respond(s, fd, ST_OK, &nv, 4);
                break;
            }
            case OP_VAR_SCALE_F32: {
                Variable* v = s->find_var(name);
                if (!v || v->dtype != 0 || plen != 4) { ok = respond(s, fd, v ? ST_BAD_REQUEST : ST_NOT_FOUND); break; }
                float f;
                memcpy(&f, payload.data(), 4);
                float* w = reinterpret_cast<float*>(v->data.data());
                for (size_t i = 0; i < v->numel(); ++i) w[i] *= f;
                ok = respond(s, fd, ST_OK);
                break;
            }
            case OP_Q_CREATE: {
                if (plen != 4) { ok = respond(s, fd, ST_BAD_REQUEST); break; }
                uint32_t cap;
                memcpy(&cap, payload.data(), 4);
                {
                    std::unique_lock<std::shared_mutex> l(s->reg_mu);
                    if (!s->queues.count(name)) {
                        auto q = std::make_unique<Queue>();
                        q->capacity = cap ? cap : 1;
                        s->queues.emplace(name, std::move(q));
                    }
                }
                ok = respond(s, fd, ST_OK);
                break;
            }
            case OP_Q_ENQUEUE: {
                Queue* q = s->find_queue(name);
                if (!q) { ok = respond(s, fd, ST_NOT_FOUND); break; }
                if (plen < 4 || plen % 4) { ok = respond(s, fd, ST_BAD_REQUEST); break; }
                const size_t n = plen / 4;
                int32_t st = ST_OK;
                for (size_t i = 0; i < n && st == ST_OK; ++i) {
                    int32_t val;
                    memcpy(&val, payload.data() + 4 * i, 4);
                    std::unique_lock<std::mutex> l(q->mu);
                    while (q->items.size() >= q->capacity && !s->stopping.load())
                        q->not_full.wait_for(l, std::chrono::milliseconds(200));
                    if (s->stopping.load()) { st = ST_STOPPED; break; }
                    q->items.push_back(val);
                    q->not_empty.notify_one();
                }
                ok = respond(s, fd, st);
                break;
            }
            case OP_Q_DEQUEUE: {
                Queue* q = s->find_queue(name);
                if (!q) { ok = respond(s, fd, ST_NOT_FOUND); break; }
                double timeout = 0.0;
                if (plen == 8) memcpy(&timeout, payload.data(), 8);
                auto deadline = std::chrono::steady_clock::now() + std::chrono::duration<double>(timeout);
                int32_t val = 0, st = ST_OK;
                {
                    std::unique_lock<std::mutex> l(q->mu);
                    while (q->items.empty()) {
                        if (s->stopping.load()) { st = ST_STOPPED; break; }
                        if (timeout > 0 && std::chrono::steady_clock::now() >= deadline) { st = ST_TIMEOUT; break; }
                        q->not_empty.wait_for(l, std::chrono::milliseconds(100));
                    }
                    if (st == ST_OK) {
                        val = q->items.front();
                        q->items.pop_front();
                        q->not_full.notify_one();
                    }
                }
                ok = respond(s, fd, st, &val, 4);
                break;
            }
            case OP_Q_SIZE: {
                Queue* q = s->find_queue(name);
                if (!q) { ok = respond(s, fd, ST_NOT_FOUND); break; }
                uint32_t n;
                {
                    std::lock_guard<std::mutex> l(q->mu);
                    n = static_cast<uint32_t>(q->items.size());
                }
                ok = respond(s, fd, ST_OK, &n, 4);
                break;
            }
            case OP_ACC_DECLARE: {
                Variable* v = s->find_var(name);
                if (!v) { ok = respond(s, fd, ST_NOT_FOUND); break; }
                {
                    std::unique_lock<std::shared_mutex> l(s->reg_mu);
                    if (!s->accs.count(name)) {
                        auto a = std::make_unique<Accumulator>();
                        a->sum.assign(v->numel(), 0.0f);
                        s->accs.emplace(name, std::move(a));
                    }
                }
                ok = respond(s, fd, ST_OK);
                break;
            }
            case OP_ACC_APPLY: {
                Accumulator* a = s->find_acc(name);
                if (!a) { ok = respond(s, fd, ST_NOT_FOUND); break; }
                if (plen != 8 + a->sum.size() * 4) { ok = respond(s, fd, ST_SHAPE); break; }
                int64_t local_step;
                memcpy(&local_step, payload.data(), 8);
                uint32_t accepted = 0;
                {
                    std::lock_guard<std::mutex> l(a->mu);
                    if (local_step >= a->global_step) {     // stale gradients are dropped
                        std::vector<float> g(a->sum.size());
                        memcpy(g.data(), payload.data() + 8, a->sum.size() * 4);
                        for (size_t i = 0; i < g.size(); ++i) a->sum[i] += g[i];
                        a->count++;
                        accepted = 1;
                    } else {
                        a->dropped++;
                    }
                }
                a->cv.notify_all();
                ok = respond(s, fd, ST_OK, &accepted, 4);
                break;
            }
            case OP_ACC_TAKE_APPLY: {
                Accumulator* a = s->find_acc(name);
                Variable* v = s->find_var(name);
                if (!a || !v) { ok = respond(s, fd, ST_NOT_FOUND); break; }
                if (plen < 4) { ok = respond(s, fd, ST_BAD_REQUEST); break; }
                uint32_t required;
                memcpy(&required, payload.data(), 4);
                std::vector<uint8_t> rest(payload.begin() + 4, payload.end());
                ApplyHdr h;
                Variable* slots[2] = {nullptr, nullptr};
                if (!parse_apply(s, rest, &h, slots)) { ok = respond(s, fd, ST_BAD_REQUEST); break; }
                if (!v->initialized.load()) { ok = respond(s, fd, ST_UNINITIALIZED); break; }
                if (!slots_fit(h, slots, v) || a->sum.size() != v->numel()) { ok = respond(s, fd, ST_SHAPE); break; }
                uint32_t used = 0;
                int32_t st = ST_OK;
                {
                    std::unique_lock<std::mutex> l(a->mu);
                    while (a->count < required) {
                        if (s->stopping.load()) { st = ST_STOPPED; break; }
                        a->cv.wait_for(l, std::chrono::milliseconds(100));
                    }
                    if (st == ST_OK) {
                        used = a->count;
                        apply_dense(h, reinterpret_cast<float*>(v->data.data()), a->sum.data(),
                                    slots[0] ? reinterpret_cast<float*>(slots[0]->data.data()) : nullptr,
                                    slots[1] ? reinterpret_cast<float*>(slots[1]->data.data()) : nullptr,
                                    v->numel(), 1.0f / static_cast<float>(used));
                        std::fill(a->sum.begin(), a->sum.end(), 0.0f);
                        a->count = 0;
                        // the accumulator's step advances WITH the take (TF's ConditionalAccumulator does the
                        // same inside TakeGrad): a backup worker's gradient computed on the old parameters
                        // that arrives from now on is stale and dropped, not averaged into the next step
                        a->global_step += 1;
                        s->n_applies++;
                    }
                }
                ok = respond(s, fd, st, &used, 4);
                break;
            }
            case OP_ACC_SET_STEP: {
                Accumulator* a = s->find_acc(name);
                if (!a || plen != 8) { ok = respond(s, fd, a ? ST_BAD_REQUEST : ST_NOT_FOUND); break; }
                int64_t step;
                memcpy(&step, payload.data(), 8);
                {
                    std::lock_guard<std::mutex> l(a->mu);
                    if (step > a->global_step) a->global_step = step;   // the take already advanced it by one
                }
                ok = respond(s, fd, ST_OK);
                break;
            }
            case OP_STATS: {
                uint64_t st[5] = {s->n_requests.load(), s->n_applies.load(), s->bytes_in.load(), s->bytes_out.load(), 0};
                {
                    std::shared_lock<std::shared_mutex> l(s->reg_mu);
                    for (auto& kv : s->accs) st[4] += kv.second->dropped;
                }
                ok = respond(s, fd, ST_OK, st, sizeof(st));
                break;
            }
            case OP_SHUTDOWN:
                respond(s, fd, ST_OK);
                s->stopping.store(true);
                ok = false;
                break;
            default:
                ok = respond(s, fd, ST_BAD_REQUEST);
        }
        } catch (const std::exception&) {
            // bad_alloc / length_error from a hostile size: answer, keep the process alive
            ok = respond(s, fd, ST_BAD_REQUEST);
        }
        if (!ok) break;
    }
    {
        // forget the descriptor BEFORE closing it: psb_ps_stop must never shut down a number the
        // process has meanwhile reused for something else
        std::lock_guard<std::mutex> l(s->conn_mu);
        for (size_t i = 0; i < s->conn_fds.size(); ++i)
            if (s->conn_fds[i] == fd) { s->conn_fds[i] = s->conn_fds.back(); s->conn_fds.pop_back(); break; }
    }
    ::close(fd);
}

void accept_loop(Server* s) {
    while (!s->stopping.load()) {
        sockaddr_in peer{};
        socklen_t len = sizeof(peer);
        int fd = ::accept(s->listen_fd, reinterpret_cast<sockaddr*>(&peer), &len);
        if (fd < 0) {
            if (s->stopping.load()) break;
            std::this_thread::sleep_for(std::chrono::milliseconds(5));
            continue;
        }
        {
            std::lock_guard<std::mutex> l(s->conn_mu);
            s->conn_fds.push_back(fd);
        }
        std::thread(handle, s, fd).detach();
    }
}

std::mutex g_mu;
std::vector<Server*> g_servers;

}  // namespace

extern "C" {

// Starts a server bound to host:port (port 0 = pick one).  Returns a handle >= 0, or < 0.
int psb_ps_start(const char* host, int port, int* bound_port) {
    int fd = ::socket(AF_INET, SOCK_STREAM, 0);
    if (fd < 0) return -1;
    int one = 1;
    setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    sockaddr_in addr{};
    addr.sin_family = AF_INET;
    addr.sin_port = htons(static_cast<uint16_t>(port));
    addr.sin_addr.s_addr = htonl(INADDR_ANY);
    if (host && *host && strcmp(host, "0.0.0.0") != 0) {
        if (inet_pton(AF_INET, host, &addr.sin_addr) != 1) {
            if (strcmp(host, "localhost") == 0) addr.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
            else addr.sin_addr.s_addr = htonl(INADDR_ANY);      // unresolvable name: listen on all
        }
    }
    if (::bind(fd, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) != 0) { ::close(fd); return -2; }
    if (::listen(fd, 128) != 0) { ::close(fd); return -3; }
    socklen_t len = sizeof(addr);
    getsockname(fd, reinterpret_cast<sockaddr*>(&addr), &len);
    Server* s = new Server();
    s->listen_fd = fd;
    s->port = ntohs(addr.sin_port);
    if (bound_port) *bound_port = s->port;
    s->accept_thread = std::thread(accept_loop, s);
    std::lock_guard<std::mutex> l(g_mu);
    g_servers.push_back(s);
    return static_cast<int>(g_servers.size()) - 1;
}

int psb_ps_stop(int handle) {
    Server* s = nullptr;
    {
        std::lock_guard<std::mutex> l(g_mu);
        if (handle < 0 || handle >= static_cast<int>(g_servers.size()) || !g_servers[handle]) return -1;
        s = g_servers[handle];
        g_servers[handle] = nullptr;
    }
    s->stopping.store(true);
    ::shutdown(s->listen_fd, SHUT_RDWR);
    ::close(s->listen_fd);
    if (s->accept_thread.joinable()) s->accept_thread.join();
    {
        std::lock_guard<std::mutex> l(s->conn_mu);
        for (int fd : s->conn_fds) ::shutdown(fd, SHUT_RDWR);
    }
    // connection threads are detached and may still be unwinding: leak the Server on purpose
    return 0;
}

int psb_ps_port(int handle) {
    std::lock_guard<std::mutex> l(g_mu);
    if (handle < 0 || handle >= static_cast<int>(g_servers.size()) || !g_servers[handle]) return -1;
    return g_servers[handle]->port;
}

}  // extern "C"
