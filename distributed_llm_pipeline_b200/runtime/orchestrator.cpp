// dlp-orchestrator: native HTTP / SSE front-end, wire-compatible with the reference's Rust orchestrator.
//
// Parity (orchestrator/src/main.rs): POST /chat {"prompt": string} -> text/event-stream, one event per message
//   data: {"msg_type":"log"|"token","content":"..."}\n\n        (:18-27, :63-95)
// keep-alive comment every second (:97), permissive CORS + static-file fallback (:100-111), default port 3005.
// The reference spawns one llama-cli per request and forwards its stdout/stderr.  This front-end instead talks to a
// persistent engine service (distributed_llm_pipeline_b200.serving.backend) over a local TCP socket with newline-
// delimited JSON, can spawn and supervise that service (restart on exit = the "auto-heal" the design report only
// proposes, PDF p.7), bounds the number of in-flight requests, forwards the service's stderr as `log` events, and
// cancels the backend request when the browser disconnects.
//
// Build: g++ -O2 -std=c++17 -pthread orchestrator.cpp -o dlp-orchestrator
#include <arpa/inet.h>
#include <errno.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <signal.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <sys/wait.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <fstream>
#include <map>
#include <mutex>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

namespace {

struct Config {
    std::string host = "0.0.0.0";
    int port = 3005;
    std::string static_dir = "static";
    std::string backend_host = "127.0.0.1";
    int backend_port = 3006;
    std::string spawn_cmd;        // optional: command line of the engine service to supervise
    int n_predict = 200;
    int max_inflight = 64;
    int backend_wait_s = 600;     // how long /chat waits for the engine service to come up
};

Config g_cfg;
std::atomic<int> g_inflight{0};
std::atomic<bool> g_stop{false};

// ---- ring of recent backend stderr lines (the "system log" pane) ------------------------------------------
std::mutex g_log_mu;
std::deque<std::string> g_log_ring;
void log_push(const std::string& line) {
    std::lock_guard<std::mutex> lk(g_log_mu);
    g_log_ring.push_back(line);
    while (g_log_ring.size() > 200) g_log_ring.pop_front();
    fputs(line.c_str(), stderr);
}
std::vector<std::string> log_snapshot() {
    std::lock_guard<std::mutex> lk(g_log_mu);
    return std::vector<std::string>(g_log_ring.begin(), g_log_ring.end());
}

// ---- JSON helpers (strings only; the wire format needs nothing else) -------------------------------------------
std::string json_escape(const std::string& s) {
    std::string o;
    o.reserve(s.size() + 8);
    for (unsigned char c : s) {
        switch (c) {
            case '"': o += "\\\""; break;
            case '\\': o += "\\\\"; break;
            case '\n': o += "\\n"; break;
            case '\r': o += "\\r"; break;
            case '\t': o += "\\t"; break;
            case '\b': o += "\\b"; break;
            case '\f': o += "\\f"; break;
            default:
                if (c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; }
                else o += (char)c;
        }
    }
    return o;
}

void utf8_append(std::string& o, unsigned cp) {
    if (cp < 0x80) o += (char)cp;
    else if (cp < 0x800) { o += (char)(0xC0 | (cp >> 6)); o += (char)(0x80 | (cp & 0x3F)); }
    else if (cp < 0x10000) { o += (char)(0xE0 | (cp >> 12)); o += (char)(0x80 | ((cp >> 6) & 0x3F)); o += (char)(0x80 | (cp & 0x3F)); }
    else { o += (char)(0xF0 | (cp >> 18)); o += (char)(0x80 | ((cp >> 12) & 0x3F)); o += (char)(0x80 | ((cp >> 6) & 0x3F)); o += (char)(0x80 | (cp & 0x3F)); }
}

// Parses the JSON string starting at s[i] == '"'; returns false on malformed input.
bool json_parse_string(const std::string& s, size_t& i, std::string& out) {
    if (i >= s.size() || s[i] != '"') return false;
    ++i;
    out.clear();
    while (i < s.size()) {
        char c = s[i++];
        if (c == '"') return true;
        if (c != '\\') { out += c; continue; }
        if (i >= s.size()) return false;
        char e = s[i++];
        switch (e) {
            case '"': out += '"'; break;
            case '\\': out += '\\'; break;
            case '/': out += '/'; break;
            case 'n': out += '\n'; break;
            case 'r': out += '\r'; break;
            case 't': out += '\t'; break;
            case 'b': out += '\b'; break;
            case 'f': out += '\f'; break;
            case 'u': {
                if (i + 4 > s.size()) return false;
                unsigned cp = (unsigned)strtoul(s.substr(i, 4).c_str(), nullptr, 16);
                i += 4;
                if (cp >= 0xD800 && cp <= 0xDBFF && i + 6 <= s.size() && s[i] == '\\' && s[i + 1] == 'u') {
                    unsigned lo = (unsigned)strtoul(s.substr(i + 2, 4).c_str(), nullptr, 16);
                    if (lo >= 0xDC00 && lo <= 0xDFFF) { cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00); i += 6; }
                }
                utf8_append(out, cp);
                break;
            }
            default: return false;
        }
    }
    return false;
}

// Finds top-level key `key` in a flat JSON object and returns its string value.
bool json_get_string(const std::string& body, const std::string& key, std::string& out) {
    size_t i = 0;
    int depth = 0;
    while (i < body.size()) {
        char c = body[i];
        if (c == '"') {
            std::string k;
            size_t j = i;
            if (!json_parse_string(body, j, k)) return false;
            size_t p = j;
            while (p < body.size() && isspace((unsigned char)body[p])) ++p;
            if (depth == 1 && p < body.size() && body[p] == ':' && k == key) {
                ++p;
                while (p < body.size() && isspace((unsigned char)body[p])) ++p;
                return json_parse_string(body, p, out);
            }
            i = j;
            continue;
        }
        if (c == '{' || c == '[') ++depth;
        if (c == '}' || c == ']') --depth;
        ++i;
    }
    return false;
}

// ---- socket helpers --------------------------------------------------------------------------------------------
bool send_all(int fd, const char* p, size_t n) {
    while (n) {
        ssize_t w = ::send(fd, p, n, MSG_NOSIGNAL);
        if (w < 0) { if (errno == EINTR) continue; return false; }
        p += w; n -= (size_t)w;
    }
    return true;
}
bool send_all(int fd, const std::string& s) { return send_all(fd, s.data(), s.size()); }

bool send_chunk(int fd, const std::string& data) {  // HTTP/1.1 chunked transfer encoding (what hyper emits for SSE)
    char head[32];
    int n = snprintf(head, sizeof head, "%zx\r\n", data.size());
    return send_all(fd, head, (size_t)n) && send_all(fd, data) && send_all(fd, "\r\n", 2);
}

int connect_to(const std::string& host, int port) {
    addrinfo hints{}, *res = nullptr;
    hints.ai_family = AF_UNSPEC;
    hints.ai_socktype = SOCK_STREAM;
    if (getaddrinfo(host.c_str(), std::to_string(port).c_str(), &hints, &res) != 0) return -1;
    int fd = -1;
    for (addrinfo* a = res; a; a = a->ai_next) {
        fd = ::socket(a->ai_family, a->ai_socktype, a->ai_protocol);
        if (fd < 0) continue;
        if (::connect(fd, a->ai_addr, a->ai_addrlen) == 0) break;
        ::close(fd);
        fd = -1;
    }
    freeaddrinfo(res);
    if (fd >= 0) { int one = 1; setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one); }
    return fd;
}

const char* CORS_HEADERS =
    "access-control-allow-origin: *\r\n"
    "vary: origin, access-control-request-method, access-control-request-headers\r\n";

std::string sse_event(const std::string& type, const std::string& content) {
    return "data: {\"msg_type\":\"" + type + "\",\"content\":\"" + json_escape(content) + "\"}\n\n";
}

// ---- HTTP request parsing ----------------------------------------------------------------------------------------
struct HttpRequest {
    std::string method, path, target;
    std::map<std::string, std::string> headers;
    std::string body;
};

bool read_request(int fd, HttpRequest& rq) {
    std::string buf;
    char tmp[4096];
    size_t hdr_end = std::string::npos;
    while (hdr_end == std::string::npos) {
        ssize_t n = ::recv(fd, tmp, sizeof tmp, 0);
        if (n <= 0) return false;
        buf.append(tmp, (size_t)n);
        hdr_end = buf.find("\r\n\r\n");
        if (buf.size() > (1u << 20)) return false;
    }
    std::istringstream hs(buf.substr(0, hdr_end));
    std::string line, target, version;
    if (!std::getline(hs, line)) return false;
    std::istringstream rl(line);
    rl >> rq.method >> target >> version;
    rq.target = target;
    rq.path = target.substr(0, target.find('?'));
    while (std::getline(hs, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        size_t c = line.find(':');
        if (c == std::string::npos) continue;
        std::string k = line.substr(0, c), v = line.substr(c + 1);
        for (auto& ch : k) ch = (char)tolower((unsigned char)ch);
        while (!v.empty() && v.front() == ' ') v.erase(v.begin());
        rq.headers[k] = v;
    }
    size_t want = 0;
    auto it = rq.headers.find("content-length");
    if (it != rq.headers.end()) want = (size_t)strtoull(it->second.c_str(), nullptr, 10);
    if (want > (64u << 20)) return false;
    rq.body = buf.substr(hdr_end + 4);
    while (rq.body.size() < want) {
        ssize_t n = ::recv(fd, tmp, sizeof tmp, 0);
        if (n <= 0) return false;
        rq.body.append(tmp, (size_t)n);
    }
    rq.body.resize(want);
    return true;
}

void send_simple(int fd, const char* status, const std::string& ctype, const std::string& body, bool head_only = false) {
    std::string h = std::string("HTTP/1.1 ") + status + "\r\ncontent-type: " + ctype + "\r\ncontent-length: " + std::to_string(body.size()) +
                    "\r\n" + CORS_HEADERS + "connection: close\r\n\r\n";
    send_all(fd, h);
    if (!head_only) send_all(fd, body);
}

const char* mime_of(const std::string& p) {
    auto ends = [&](const char* e) { size_t n = strlen(e); return p.size() >= n && p.compare(p.size() - n, n, e) == 0; };
    if (ends(".html") || ends(".htm")) return "text/html; charset=utf-8";
    if (ends(".js")) return "text/javascript; charset=utf-8";
    if (ends(".css")) return "text/css; charset=utf-8";
    if (ends(".json")) return "application/json";
    if (ends(".svg")) return "image/svg+xml";
    if (ends(".png")) return "image/png";
    if (ends(".ico")) return "image/x-icon";
    if (ends(".txt")) return "text/plain; charset=utf-8";
    return "application/octet-stream";
}

void serve_static(int fd, const HttpRequest& rq) {
    std::string rel = rq.path;
    if (rel.find("..") != std::string::npos) { send_simple(fd, "404 Not Found", "text/plain", ""); return; }
    if (rel.empty() || rel.back() == '/') rel += "index.html";
    std::string full = g_cfg.static_dir + rel;
    struct stat st{};
    if (stat(full.c_str(), &st) == 0 && S_ISDIR(st.st_mode)) full += "/index.html";
    std::ifstream f(full, std::ios::binary);
    if (!f) { send_simple(fd, "404 Not Found", "text/plain", ""); return; }
    std::stringstream ss;
    ss << f.rdbuf();
    send_simple(fd, "200 OK", mime_of(full), ss.str(), rq.method == "HEAD");
}

// ---- /chat ---------------------------------------------------------------------------------------------------------
bool client_gone(int fd) {
    pollfd p{fd, POLLIN | POLLRDHUP, 0};
    if (poll(&p, 1, 0) > 0) {
        if (p.revents & (POLLRDHUP | POLLHUP | POLLERR)) return true;
        char c;
        if ((p.revents & POLLIN) && ::recv(fd, &c, 1, MSG_PEEK | MSG_DONTWAIT) == 0) return true;
    }
    return false;
}

void handle_chat(int fd, const HttpRequest& rq) {
    std::string prompt;
    if (!json_get_string(rq.body, "prompt", prompt)) {
        send_simple(fd, "422 Unprocessable Entity", "text/plain; charset=utf-8", "Failed to deserialize the JSON body into the target type");
        return;
    }
    if (g_inflight.load() >= g_cfg.max_inflight) {
        send_simple(fd, "503 Service Unavailable", "text/plain; charset=utf-8", "too many requests in flight");
        return;
    }
    struct Guard { Guard() { ++g_inflight; } ~Guard() { --g_inflight; } } guard;
    std::string head = std::string("HTTP/1.1 200 OK\r\ncontent-type: text/event-stream\r\ncache-control: no-cache\r\n") + CORS_HEADERS +
                       "transfer-encoding: chunked\r\n\r\n";
    if (!send_all(fd, head)) return;
    for (const auto& l : log_snapshot())
        if (!send_chunk(fd, sse_event("log", l))) return;

    // connect to the engine service (it may still be loading the model): keep the stream alive meanwhile
    int be = -1;
    auto t0 = std::chrono::steady_clock::now();
    while ((be = connect_to(g_cfg.backend_host, g_cfg.backend_port)) < 0) {
        if (std::chrono::steady_clock::now() - t0 > std::chrono::seconds(g_cfg.backend_wait_s) || g_stop) {
            send_chunk(fd, sse_event("log", "error: engine service unreachable\n"));
            send_all(fd, "0\r\n\r\n", 5);
            return;
        }
        if (!send_chunk(fd, ":\n\n") || client_gone(fd)) return;
        std::this_thread::sleep_for(std::chrono::seconds(1));
    }
    std::string req = "{\"op\":\"generate\",\"prompt\":\"" + json_escape(prompt) + "\",\"n_predict\":" + std::to_string(g_cfg.n_predict) + "}\n";
    if (!send_all(be, req)) { ::close(be); send_chunk(fd, sse_event("log", "error: engine service closed the connection\n")); send_all(fd, "0\r\n\r\n", 5); return; }

    std::string buf;
    char tmp[8192];
    bool done = false;
    while (!done) {
        pollfd p{be, POLLIN, 0};
        int r = poll(&p, 1, 1000);               // 1 s keep-alive interval, as the reference
        if (r == 0) {
            if (!send_chunk(fd, ":\n\n") || client_gone(fd)) break;   // client left: closing `be` cancels the request
            continue;
        }
        if (r < 0) { if (errno == EINTR) continue; break; }
        ssize_t n = ::recv(be, tmp, sizeof tmp, 0);
        if (n <= 0) {
            send_chunk(fd, sse_event("log", "error: engine service went away (it will be restarted if supervised)\n"));
            break;
        }
        buf.append(tmp, (size_t)n);
        size_t nl;
        while ((nl = buf.find('\n')) != std::string::npos) {
            std::string line = buf.substr(0, nl);
            buf.erase(0, nl + 1);
            std::string t, c;
            if (!json_get_string(line, "t", t)) continue;
            json_get_string(line, "c", c);
            if (t == "done") { done = true; break; }
            std::string ev = sse_event(t == "token" ? "token" : "log", t == "error" ? "error: " + c + "\n" : c);
            if (!send_chunk(fd, ev)) { done = true; break; }
        }
    }
    ::close(be);
    send_all(fd, "0\r\n\r\n", 5);
}

void proxy_get(int fd, const std::string& path) {  // /health, /metrics, /slots of the engine service
    int be = connect_to(g_cfg.backend_host, g_cfg.backend_port);
    if (be < 0) { send_simple(fd, "503 Service Unavailable", "application/json", "{\"status\":\"loading\"}"); return; }
    std::string req = "{\"op\":\"get\",\"path\":\"" + json_escape(path) + "\"}\n";
    send_all(be, req);
    std::string buf;
    char tmp[8192];
    ssize_t n;
    while ((n = ::recv(be, tmp, sizeof tmp, 0)) > 0) buf.append(tmp, (size_t)n);
    ::close(be);
    std::string c;
    if (!json_get_string(buf, "c", c)) c = "{}";
    send_simple(fd, "200 OK", path == "/metrics" ? "text/plain; version=0.0.4" : "application/json", c);
}

// Every API route the front-end does not serve itself (/completion, /v1/*, /embedding, /tokenize, /slots/.., /props, ...) is
// tunnelled to the engine service: request line + headers as one JSON line, the body as raw bytes, and the service's complete
// HTTP response (chunked SSE streams included) relayed verbatim.  Closing either side cancels the request.
void proxy_http(int fd, const HttpRequest& rq) {
    int be = connect_to(g_cfg.backend_host, g_cfg.backend_port);
    if (be < 0) { send_simple(fd, "503 Service Unavailable", "application/json", "{\"error\":\"engine service is loading\"}"); return; }
    std::string hdr = "{\"op\":\"http\",\"method\":\"" + json_escape(rq.method) + "\",\"target\":\"" + json_escape(rq.target) + "\",\"headers\":{";
    bool first = true;
    for (const auto& kv : rq.headers) {
        if (kv.first != "authorization" && kv.first != "x-api-key" && kv.first != "content-type" && kv.first != "accept") continue;
        hdr += std::string(first ? "" : ",") + "\"" + json_escape(kv.first) + "\":\"" + json_escape(kv.second) + "\"";
        first = false;
    }
    hdr += "},\"body_len\":" + std::to_string(rq.body.size()) + "}\n";
    if (!send_all(be, hdr) || !send_all(be, rq.body)) { ::close(be); send_simple(fd, "502 Bad Gateway", "application/json", "{\"error\":\"engine service closed the connection\"}"); return; }
    char tmp[16384];
    for (;;) {
        pollfd p{be, POLLIN, 0};
        int r = poll(&p, 1, 1000);
        if (r == 0) { if (client_gone(fd)) break; continue; }
        if (r < 0) { if (errno == EINTR) continue; break; }
        ssize_t n = ::recv(be, tmp, sizeof tmp, 0);
        if (n <= 0) break;
        if (!send_all(fd, tmp, (size_t)n)) break;
    }
    ::close(be);
}

void handle_connection(int fd) {
    int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
    HttpRequest rq;
    if (read_request(fd, rq)) {
        if (rq.method == "OPTIONS") {
            std::string h = std::string("HTTP/1.1 200 OK\r\n") + CORS_HEADERS +
                            "access-control-allow-methods: *\r\naccess-control-allow-headers: *\r\ncontent-length: 0\r\nconnection: close\r\n\r\n";
            send_all(fd, h);
        } else if (rq.method == "POST" && rq.path == "/chat") {
            handle_chat(fd, rq);
        } else if (rq.method == "GET" && (rq.path == "/health" || rq.path == "/v1/health" || rq.path == "/metrics" || rq.path == "/slots")) {
            proxy_get(fd, rq.path);
        } else if (rq.method == "POST" || (rq.method == "GET" && (rq.path == "/props" || rq.path == "/v1/models" || rq.path == "/models" ||
                                                                   rq.path == "/lora-adapters" || rq.path == "/api/tags"))) {
            proxy_http(fd, rq);
        } else if (rq.method == "GET" || rq.method == "HEAD") {
            serve_static(fd, rq);
        } else {
            send_simple(fd, "405 Method Not Allowed", "text/plain", "");
        }
    }
    ::shutdown(fd, SHUT_RDWR);
    ::close(fd);
}

// ---- engine-service supervision ---------------------------------------------------------------------------------------
void supervise() {
    int restarts = 0;
    while (!g_stop) {
        int ep[2];
        if (pipe(ep) != 0) return;
        pid_t pid = fork();
        if (pid == 0) {
            dup2(ep[1], 2);
            ::close(ep[0]);
            ::close(ep[1]);
            execl("/bin/sh", "sh", "-c", g_cfg.spawn_cmd.c_str(), (char*)nullptr);
            _exit(127);
        }
        ::close(ep[1]);
        log_push("orchestrator: engine service started (pid " + std::to_string(pid) + ", restart #" + std::to_string(restarts) + ")\n");
        std::string acc;
        char tmp[1024];
        ssize_t n;
        while ((n = ::read(ep[0], tmp, sizeof tmp)) > 0) {   // 1024-byte stderr chunks, like the reference (:63-80)
            acc.append(tmp, (size_t)n);
            size_t nl;
            while ((nl = acc.find('\n')) != std::string::npos) { log_push(acc.substr(0, nl + 1)); acc.erase(0, nl + 1); }
        }
        ::close(ep[0]);
        int status = 0;
        waitpid(pid, &status, 0);
        if (g_stop) break;
        ++restarts;
        log_push("orchestrator: engine service exited (status " + std::to_string(status) + "), restarting in 1 s\n");
        std::this_thread::sleep_for(std::chrono::seconds(1));
    }
}

void usage() {
    fprintf(stderr,
            "dlp-orchestrator [--host H] [--port 3005] [--static DIR] [--backend HOST:PORT] [--spawn 'CMD'] [-n N_PREDICT] [--max-inflight N]\n"
            "  env: DLP_HOST DLP_PORT DLP_STATIC DLP_BACKEND DLP_SPAWN DLP_N_PREDICT\n");
}

}  // namespace

int main(int argc, char** argv) {
    signal(SIGPIPE, SIG_IGN);
    auto env = [](const char* k, const char* d) { const char* v = getenv(k); return std::string(v ? v : d); };
    g_cfg.host = env("DLP_HOST", "0.0.0.0");
    g_cfg.port = atoi(env("DLP_PORT", "3005").c_str());
    g_cfg.static_dir = env("DLP_STATIC", "static");
    g_cfg.spawn_cmd = env("DLP_SPAWN", "");
    g_cfg.n_predict = atoi(env("DLP_N_PREDICT", "200").c_str());
    std::string backend = env("DLP_BACKEND", "127.0.0.1:3006");
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto next = [&]() -> std::string { if (i + 1 >= argc) { usage(); exit(2); } return argv[++i]; };
        if (a == "--host") g_cfg.host = next();
        else if (a == "--port") g_cfg.port = atoi(next().c_str());
        else if (a == "--static") g_cfg.static_dir = next();
        else if (a == "--backend") backend = next();
        else if (a == "--spawn") g_cfg.spawn_cmd = next();
        else if (a == "-n" || a == "--predict") g_cfg.n_predict = atoi(next().c_str());
        else if (a == "--max-inflight") g_cfg.max_inflight = atoi(next().c_str());
        else if (a == "--backend-wait") g_cfg.backend_wait_s = atoi(next().c_str());
        else { usage(); return a == "-h" || a == "--help" ? 0 : 2; }
    }
    size_t c = backend.rfind(':');
    if (c != std::string::npos) { g_cfg.backend_host = backend.substr(0, c); g_cfg.backend_port = atoi(backend.substr(c + 1).c_str()); }

    int ls = ::socket(AF_INET, SOCK_STREAM, 0);
    int one = 1;
    setsockopt(ls, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    sockaddr_in addr{};
    addr.sin_family = AF_INET;
    addr.sin_port = htons((uint16_t)g_cfg.port);
    if (inet_pton(AF_INET, g_cfg.host.c_str(), &addr.sin_addr) != 1) addr.sin_addr.s_addr = INADDR_ANY;
    if (bind(ls, (sockaddr*)&addr, sizeof addr) != 0 || listen(ls, 128) != 0) { perror("bind/listen"); return 1; }
    socklen_t alen = sizeof addr;
    getsockname(ls, (sockaddr*)&addr, &alen);
    printf("dlp-orchestrator ready: http://localhost:%d (engine service %s:%d)\n", ntohs(addr.sin_port), g_cfg.backend_host.c_str(), g_cfg.backend_port);
    fflush(stdout);
    std::thread sup;
    if (!g_cfg.spawn_cmd.empty()) sup = std::thread(supervise);
    while (!g_stop) {
        int fd = ::accept(ls, nullptr, nullptr);
        if (fd < 0) { if (errno == EINTR) continue; break; }
        std::thread(handle_connection, fd).detach();
    }
    g_stop = true;
    if (sup.joinable()) sup.detach();
    return 0;
}
