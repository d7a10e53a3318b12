// service.cpp — MockService / Server mirrors (see service.hpp for the Go ↔ C++ map).
#include "service.hpp"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <ctime>

namespace polykey {

namespace {
void json_string(const std::string& s, std::string& out)
{
    out.push_back('"');
    for (unsigned char c : s) {
        switch (c) {
        case '"':  out += "\\\""; break;
        case '\\': out += "\\\\"; break;
        case '\n': out += "\\n"; break;
        case '\r': out += "\\r"; break;
        case '\t': out += "\\t"; break;
        default:
            if (c < 0x20) { char b[8]; std::snprintf(b, sizeof b, "\\u%04x", c); out += b; }
            else out.push_back((char)c);
        }
    }
    out.push_back('"');
}

void json_value(const Value& v, std::string& out);

void json_struct(const Struct& s, std::string& out)
{
    out.push_back('{');
    bool first = true;
    for (const auto& kv : s) {
        if (!first) out.push_back(',');
        first = false;
        json_string(kv.first, out);
        out.push_back(':');
        json_value(kv.second, out);
    }
    out.push_back('}');
}

void json_value(const Value& v, std::string& out)
{
    switch (v.v.index()) {
    case 0: out += "null"; break;
    case 1: out += std::get<bool>(v.v) ? "true" : "false"; break;
    case 2: {
        const double d = std::get<double>(v.v);
        char b[40];
        if (std::floor(d) == d && std::fabs(d) < 9.0e15) std::snprintf(b, sizeof b, "%lld", (long long)d);
        else std::snprintf(b, sizeof b, "%.17g", d);
        out += b;
        break;
    }
    case 3: json_string(std::get<std::string>(v.v), out); break;
    case 4: json_struct(*std::get<std::shared_ptr<Struct>>(v.v), out); break;
    default: {
        out.push_back('[');
        bool first = true;
        for (const Value& e : *std::get<std::shared_ptr<List>>(v.v)) {
            if (!first) out.push_back(',');
            first = false;
            json_value(e, out);
        }
        out.push_back(']');
    }
    }
}
}  // namespace

std::string to_json(const Struct& s) { std::string o; json_struct(s, o); return o; }
std::string to_json(const Value& v) { std::string o; json_value(v, o); return o; }

std::string rfc3339_utc(int64_t unix_seconds)
{
    std::time_t t = (std::time_t)unix_seconds;
    std::tm tmv{};
    gmtime_r(&t, &tmv);
    char b[40];
    std::strftime(b, sizeof b, "%Y-%m-%dT%H:%M:%SZ", &tmv);
    return b;
}

int64_t LoggingInterceptor::now_ns()
{
    return std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// time.Duration.String(): "1.5s", "2.25ms", "340µs", "17ns" — the largest unit that keeps the value >= 1
std::string LoggingInterceptor::format_duration(int64_t ns)
{
    char b[48];
    if (ns < 1000) std::snprintf(b, sizeof b, "%lldns", (long long)ns);
    else {
        const char* unit = ns < 1000000 ? "µs" : ns < 1000000000 ? "ms" : "s";
        const double v = ns < 1000000 ? ns / 1e3 : ns < 1000000000 ? ns / 1e6 : ns / 1e9;
        std::snprintf(b, sizeof b, "%.9g%s", v, unit);
    }
    return b;
}

void LoggingInterceptor::log(const char* level, const char* msg, const std::string& method,
                             const std::string* duration, const char* code)
{
    if (!sink_) return;
    std::string line = std::string("{\"level\":\"") + level + "\",\"msg\":\"" + msg + "\",\"method\":\"" + method + "\"";
    if (duration) line += ",\"duration\":\"" + *duration + "\"";
    if (code) line += std::string(",\"code\":\"") + code + "\"";
    sink_->push_back(line + "}");
}

std::unique_ptr<MockService> NewMockService() { return std::make_unique<MockService>(); }

// internal/service/mock.go:22-67.  ctx, parameters, secretId and metadata are ignored, as in
// the reference; an unknown tool is NOT an error (mock.go:60-66).
Result MockService::ExecuteTool(Context& ctx, const std::string& toolName, const Struct*,
                                const std::string*, const Metadata*)
{
    const int64_t now = ctx.now_unix ? ctx.now_unix : (int64_t)std::time(nullptr);
    auto resp = std::make_unique<ExecuteToolResponse>();
    resp->status = Status{200, "Tool executed successfully"};                   // mock.go:24-29
    if (toolName == "example_tool") {                                            // mock.go:33-36
        resp->output = "Mock execution of " + toolName + " at " + rfc3339_utc(now);
    } else if (toolName == "struct_tool") {                                      // mock.go:37-51
        Struct data{{"processed", Value(true)}, {"count", Value(42)}};
        resp->output = Struct{{"result", Value("success")}, {"timestamp", Value(now)}, {"data", Value(data)}};
    } else if (toolName == "file_tool") {                                        // mock.go:52-59
        resp->output = File{"example.txt", "text/plain", "This is mock file content"};
    } else {                                                                     // mock.go:60-64
        resp->output = "Unknown tool: " + toolName;
    }
    return Result{std::move(resp), std::nullopt};                                // mock.go:66
}

// internal/server/server.go:27-43
Result Server::ExecuteTool(Context& ctx, const ExecuteToolRequest& req)
{
    if (log_) {
        log_->push_back(std::string("ExecuteTool called tool_name=") + req.tool_name +
                        " has_parameters=" + (req.parameters ? "true" : "false") +
                        " has_secret_id=" + (req.secret_id ? "true" : "false") +
                        " has_metadata=" + (req.metadata ? "true" : "false"));   // server.go:28-33
    }
    Result r = service_->ExecuteTool(ctx, req.tool_name, req.parameters ? &*req.parameters : nullptr,
                                     req.secret_id ? &*req.secret_id : nullptr,
                                     req.metadata ? &*req.metadata : nullptr);   // server.go:36
    if (r.error) {
        if (log_) log_->push_back("Service ExecuteTool failed error=" + *r.error);   // server.go:38
        return Result{nullptr, r.error};                                         // server.go:39: verbatim
    }
    return r;
}

}  // namespace polykey
