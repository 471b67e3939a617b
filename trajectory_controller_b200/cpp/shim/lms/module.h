// HEADLESS SHIM of the LMS core surface the trajectory_point_controller module touches.
// The real `lmscore` is not vendored in the reference (README.md:9-13 lists it as an external dependency), so this
// file reconstructs, from the module's call sites only, what is needed to compile and drive the module without
// the LMS runtime: Module base class with the four life-cycle virtuals (include/trajectory_point_follower.h:28-30,45),
// typed data channels (src/trajectory_point_follower.cpp:9-13), services (:35), config (:64,71,92-95,292-298) and
// a logger (:80,100-103).  A build against the real LMS headers simply does not put shim/ on the include path.
#pragma once
#include <cstdint>
#include <iostream>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <typeindex>
#include <vector>

namespace lms {

// ---- time (only what the module names: Time::ZERO, now(), since(), fromMillis()) -------------------------------
class Time {
    int64_t us_;
public:
    explicit Time(int64_t us = 0) : us_(us) {}
    static const Time ZERO;
    static Time now();
    static Time fromMillis(int64_t ms) { return Time(ms * 1000); }
    Time since() const { return Time(now().us_ - us_); }
    bool operator>(const Time& o) const { return us_ > o.us_; }
    int64_t micros() const { return us_; }
};

// ---- logging --------------------------------------------------------------------------------------------------------
class LogLine {
    std::ostringstream os_;
    const char* level_;
    bool live_;
public:
    LogLine(const char* level, const std::string& tag, bool live) : level_(level), live_(live) { os_ << tag << ": "; }
    LogLine(LogLine&& o) : os_(std::move(o.os_)), level_(o.level_), live_(o.live_) { o.live_ = false; }
    ~LogLine();
    template <typename T> LogLine& operator<<(const T& v) { if (live_) os_ << v; return *this; }
};
class Logger {
public:
    int verbosity = 1;                 // 0: silent, 1: warn+error, 2: everything
    std::vector<std::string> errors;   // kept so tests can assert on error reports
    LogLine debug(const std::string& tag = "") { return LogLine("debug", tag, verbosity >= 2); }
    LogLine info(const std::string& tag = "") { return LogLine("info", tag, verbosity >= 2); }
    LogLine warn(const std::string& tag = "") { return LogLine("warn", tag, verbosity >= 1); }
    LogLine error(const std::string& tag = "") { errors.push_back(tag); return LogLine("error", tag, verbosity >= 1); }
    void time(const std::string&) {}
    void timeEnd(const std::string&) {}
};

// ---- config ----------------------------------------------------------------------------------------------------------
class Config {
    std::map<std::string, std::string> kv_;
    template <typename T> static T parse(const std::string& s) { std::istringstream is(s); T v{}; is >> v; return v; }
public:
    template <typename T> void set(const std::string& key, const T& v) { std::ostringstream os; os.precision(17); os << v; kv_[key] = os.str(); }
    template <typename T> void setArray(const std::string& key, const std::vector<T>& v)
    {
        std::ostringstream os; os.precision(9);
        for (size_t i = 0; i < v.size(); ++i) os << (i ? "," : "") << v[i];
        kv_[key] = os.str();
    }
    bool hasKey(const std::string& key) const { return kv_.count(key) != 0; }
    template <typename T> T get(const std::string& key, const T& def = T()) const
    {
        auto it = kv_.find(key);
        return it == kv_.end() ? def : parse<T>(it->second);
    }
    template <typename T> std::vector<T> getArray(const std::string& key) const
    {
        std::vector<T> out;
        auto it = kv_.find(key);
        if (it == kv_.end()) return out;
        std::istringstream is(it->second); std::string tok;
        while (std::getline(is, tok, ',')) if (!tok.empty()) out.push_back(parse<T>(tok));
        return out;
    }
};
template <> inline std::string Config::parse<std::string>(const std::string& s) { return s; }

// ---- blackboard: channels and services by name ----------------------------------------------------------------------
class Blackboard {
    struct Slot { std::shared_ptr<void> obj; std::type_index type = std::type_index(typeid(void)); };
    std::map<std::string, Slot> channels_, services_;
public:
    static Blackboard& instance();
    void clear() { channels_.clear(); services_.clear(); }
    template <typename T> std::shared_ptr<T> channel(const std::string& name)
    {
        Slot& s = channels_[name];
        if (!s.obj) { s.obj = std::make_shared<T>(); s.type = std::type_index(typeid(T)); }
        return std::static_pointer_cast<T>(s.obj);
    }
    template <typename T> void provideService(const std::string& name, std::shared_ptr<T> svc)
    {
        services_[name] = Slot{svc, std::type_index(typeid(T))};
    }
    template <typename T> std::shared_ptr<T> service(const std::string& name)
    {
        auto it = services_.find(name);
        return it == services_.end() ? std::shared_ptr<T>() : std::static_pointer_cast<T>(it->second.obj);
    }
};

template <typename T>
class ReadDataChannel {
    std::shared_ptr<T> p_;
public:
    ReadDataChannel() {}
    explicit ReadDataChannel(std::shared_ptr<T> p) : p_(p) {}
    const T* operator->() const { return p_.get(); }
    const T& operator*() const { return *p_; }
    const T* get() const { return p_.get(); }
    explicit operator bool() const { return (bool)p_; }
};
template <typename T>
class WriteDataChannel {
    std::shared_ptr<T> p_;
public:
    WriteDataChannel() {}
    explicit WriteDataChannel(std::shared_ptr<T> p) : p_(p) {}
    T* operator->() const { return p_.get(); }
    T& operator*() const { return *p_; }
    T* get() const { return p_.get(); }
    explicit operator bool() const { return (bool)p_; }
};
template <typename T>
class ServiceHandle {
    std::shared_ptr<T> p_;
public:
    explicit ServiceHandle(std::shared_ptr<T> p) : p_(p) {}
    T* operator->() const { return p_.get(); }
    bool isValid() const { return (bool)p_; }
    explicit operator bool() const { return (bool)p_; }
};

// ---- module base class ------------------------------------------------------------------------------------------------
class Module {
public:
    virtual ~Module() {}
    virtual bool initialize() = 0;
    virtual bool deinitialize() = 0;
    virtual bool cycle() = 0;
    virtual void configsChanged() {}
    Config& config() { return config_; }
    const Config& config() const { return config_; }
    Logger logger;

protected:
    template <typename T> ReadDataChannel<T> readChannel(const std::string& name)
    {
        return ReadDataChannel<T>(Blackboard::instance().channel<T>(name));
    }
    template <typename T> WriteDataChannel<T> writeChannel(const std::string& name)
    {
        return WriteDataChannel<T>(Blackboard::instance().channel<T>(name));
    }
    template <typename T> ServiceHandle<T> getService(const std::string& name)
    {
        return ServiceHandle<T>(Blackboard::instance().service<T>(name));
    }

private:
    Config config_;
};

}  // namespace lms

// The loader contract: a C factory the LMS runtime dlsym()s from the MODULE library (reference src/interface.cpp:3).
#define LMS_MODULE_INTERFACE(CLASS)                                                   \
    extern "C" {                                                                      \
    uint32_t getLmsVersion() { return 0x00010000u; }                                  \
    void* getInstance() { return static_cast<lms::Module*>(new CLASS()); }            \
    }
