// service.hpp — C++ mirror of the reference's plugin seam and message shapes.
//
// The reference is Go; there is no Go toolchain in this environment, so the host side above
// the C ABI is written in C++ with the SAME names, argument meaning and error behaviour:
//
//   Go (reference)                                               here
//   -----------------------------------------------------------  -------------------------
//   service.Service interface  (internal/service/service.go:13)  polykey::Service
//   service.MockService        (internal/service/mock.go:14-67)  polykey::MockService
//   server.Server / NewServer  (internal/server/server.go:11-43) polykey::Server
//   (resp, error) return pair                                    polykey::Result {resp, error}
//   *structpb.Struct parameters                                  polykey::Struct (JSON-like tree)
//   *string secretId                                             const std::string* (nullable)
//   *cmn.Metadata{Fields map[string]string}                      polykey::Metadata
//   pk.ExecuteToolResponse{Status, Output oneof}                 polykey::ExecuteToolResponse
//
// The message field shapes are inferred from the call sites (server.go:29-36, mock.go:24-63,
// cmd/dev_client/main.go:247-258): the .proto is not in the reference tree.
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <optional>
#include <string>
#include <variant>
#include <vector>

#include "../../include/pk_route.h"

namespace polykey {

// ---- structpb.Struct / Value -----------------------------------------------------------
struct Value;
using Struct = std::map<std::string, Value>;
using List = std::vector<Value>;
struct Value {
    std::variant<std::nullptr_t, bool, double, std::string, std::shared_ptr<Struct>, std::shared_ptr<List>> v;
    Value() : v(nullptr) {}
    Value(bool b) : v(b) {}
    Value(double d) : v(d) {}
    Value(int i) : v((double)i) {}
    Value(int64_t i) : v((double)i) {}
    Value(const char* s) : v(std::string(s)) {}
    Value(std::string s) : v(std::move(s)) {}
    Value(Struct s) : v(std::make_shared<Struct>(std::move(s))) {}
    Value(List l) : v(std::make_shared<List>(std::move(l))) {}
};

// Canonical JSON rendering (keys in map order = sorted; numbers as integers when integral).
// This is the byte payload the routing stage scans for a request's `parameters`.
std::string to_json(const Struct& s);
std::string to_json(const Value& v);

// ---- common/v2 messages ------------------------------------------------------------------
struct Metadata { std::map<std::string, std::string> fields; };
struct Status { int32_t code = 0; std::string message; };
struct File { std::string file_name, mime_type, content; };

// ---- polykey/v2 ExecuteToolResponse --------------------------------------------------------
struct ExecuteToolResponse {
    Status status;
    // oneof output: string_output | struct_output | file_output  (mock.go:33-63)
    std::variant<std::monostate, std::string, Struct, File> output;
    // Added by the batching decorator; absent when MockService is called directly.
    std::optional<pk_decision> decision;
};

struct ExecuteToolRequest {
    std::string tool_name;
    std::optional<Struct> parameters;
    std::optional<std::string> secret_id;
    std::optional<Metadata> metadata;
};

// Go's `error`: empty optional = nil.
using Error = std::optional<std::string>;
struct Result {
    std::unique_ptr<ExecuteToolResponse> resp;     // nullptr when error is set (Go: nil, err)
    Error error;
};

struct Context {
    int64_t now_unix = 0;      // replaces time.Now() (mock.go:35,40) so tests are reproducible; 0 = wall clock
    bool cancelled = false;
};

// service.Service — internal/service/service.go:13-15
class Service {
public:
    virtual ~Service() = default;
    virtual Result ExecuteTool(Context& ctx, const std::string& toolName, const Struct* parameters,
                               const std::string* secretId, const Metadata* metadata) = 0;
};

// service.MockService — internal/service/mock.go:14-67.  Stateless, concurrency-safe.
class MockService final : public Service {
public:
    Result ExecuteTool(Context& ctx, const std::string& toolName, const Struct* parameters,
                       const std::string* secretId, const Metadata* metadata) override;
};
std::unique_ptr<MockService> NewMockService();     // mock.go:17-19

// server.Server — internal/server/server.go:11-43: one log line, delegate, error pass-through.
class Server {
public:
    explicit Server(Service* service, std::vector<std::string>* log_sink = nullptr)
        : service_(service), log_(log_sink) {}
    Result ExecuteTool(Context& ctx, const ExecuteToolRequest& req);
private:
    Service* service_;
    std::vector<std::string>* log_;
};
inline Server NewServer(Service* s) { return Server(s); }   // server.go:19-24

// loggingInterceptor — cmd/polykey/main.go:25-52.  Wraps one unary call: health checks pass
// through unlogged (:29-31); otherwise "gRPC call received" before (:33) and "gRPC call finished"
// with duration and status code after (:44-48) — level ERROR and code "Unknown" when the handler
// returned a plain error (status.Code of a non-status error), "OK" otherwise.  The response and
// the error are returned untouched (:50).  Log records go to `sink` as the reference's JSON lines
// minus the timestamp.
class LoggingInterceptor {
public:
    explicit LoggingInterceptor(std::vector<std::string>* sink) : sink_(sink) {}
    template <typename Handler>
    Result Intercept(const std::string& fullMethod, Handler&& handler)
    {
        if (fullMethod == "/grpc.health.v1.Health/Check") return handler();
        const int64_t t0 = now_ns();
        log("INFO", "gRPC call received", fullMethod, nullptr, nullptr);
        Result r = handler();
        const std::string dur = format_duration(now_ns() - t0);
        const char* code = r.error ? "Unknown" : "OK";
        log(r.error ? "ERROR" : "INFO", "gRPC call finished", fullMethod, &dur, code);
        return r;
    }
    static std::string format_duration(int64_t ns);      // Go's time.Duration.String() for sub-minute spans
private:
    static int64_t now_ns();
    void log(const char* level, const char* msg, const std::string& method, const std::string* duration, const char* code);
    std::vector<std::string>* sink_;
};

std::string rfc3339_utc(int64_t unix_seconds);

}  // namespace polykey
