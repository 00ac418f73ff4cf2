// Pothos/Framework.hpp -- a header-only stand-in for the slice of the Pothos framework that the four upstream
// LoRa blocks touch, so that LoRa-SDR/LoRaDecoder.cpp, LoRaEncoder.cpp, LoRaDemod.cpp and LoRaMod.cpp compile
// UNMODIFIED from /root/reference into oracle/_ref/liblora_pothos_ref.so.
//
// TEST INFRASTRUCTURE ONLY.  It contains no LoRa arithmetic: ports backed by plain memory, packets, labels and
// signals that are merely recorded, and a registry that does nothing.  pothos_doors.cpp plays the scheduler
// (feeds the ports, calls activate() once and work() repeatedly, collects what the block posted).
//
// Surface covered (everything the four files name):
//   Pothos::Block { registerCall, registerSignal, emitSignal, setupInput, setupOutput, input(), output(),
//                   activate(), work(), getInputBufferManager(), getOutputBufferManager() }
//   Pothos::InputPort { hasMessage, popMessage, elements, buffer, consume, setReserve }
//   Pothos::OutputPort { buffer, produce, postMessage, postLabel, setReserve }
//   Pothos::BufferChunk, DType, Packet, Label, Object, Exception, InvalidArgumentException,
//   BufferManager, BufferManagerArgs, BlockRegistry, POTHOS_FCN_TUPLE
#pragma once
#include <complex>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <typeinfo>
#include <utility>
#include <vector>

#define POTHOS_FCN_TUPLE(cls, fn) #fn, &cls::fn

namespace Pothos {

class Exception : public std::runtime_error
{
public:
    Exception(const std::string& where, const std::string& what = "") : std::runtime_error(where + ": " + what) {}
};
class InvalidArgumentException : public Exception
{
public:
    InvalidArgumentException(const std::string& where, const std::string& what = "") : Exception(where, what) {}
};

// Element type of a buffer: only its size matters here.
class DType
{
public:
    DType() : _size(1) {}
    DType(const std::type_info& t) : _size(sizeOf(t)) {}
    size_t size() const { return _size; }

private:
    static size_t sizeOf(const std::type_info& t)
    {
        if (t == typeid(uint8_t) || t == typeid(int8_t) || t == typeid(char)) return 1;
        if (t == typeid(uint16_t) || t == typeid(int16_t)) return 2;
        if (t == typeid(uint32_t) || t == typeid(int32_t) || t == typeid(float)) return 4;
        if (t == typeid(std::complex<float>) || t == typeid(double) || t == typeid(uint64_t) || t == typeid(int64_t)) return 8;
        if (t == typeid(std::complex<double>)) return 16;
        throw Exception("Pothos stub DType", std::string("unknown element type ") + t.name());
    }
    size_t _size;
};

// A reference-counted slab, or a view on memory the scheduler owns.  Copies share the slab, as in Pothos.
class BufferChunk
{
public:
    BufferChunk() : address(0), length(0) {}
    BufferChunk(const DType& t, size_t numElements) : address(0), length(numElements * t.size()), dtype(t)
    {
        _store = std::make_shared<std::vector<char>>(length ? length : 1);
        address = reinterpret_cast<size_t>(_store->data());
    }
    static BufferChunk view(const DType& t, void* p, size_t numElements)
    {
        BufferChunk b;
        b.dtype = t;
        b.address = reinterpret_cast<size_t>(p);
        b.length = numElements * t.size();
        return b;
    }
    template <typename T> T as() const { return reinterpret_cast<T>(address); }
    size_t elements() const { return length / dtype.size(); }

    size_t address;
    size_t length; // bytes
    DType dtype;

private:
    std::shared_ptr<std::vector<char>> _store;
};

// A type-erased value: messages on ports and label payloads.
class Object
{
public:
    Object() {}
    template <typename T> Object(const T& v) : _p(std::make_shared<T>(v)), _t(&typeid(T)) {}
    template <typename T> const T& extract() const
    {
        if (!_p || *_t != typeid(T)) throw Exception("Pothos stub Object::extract", "type mismatch");
        return *static_cast<const T*>(_p.get());
    }
    bool empty() const { return !_p; }

private:
    std::shared_ptr<void> _p;
    const std::type_info* _t = nullptr;
};

struct Label
{
    Label() : index(0) {}
    Label(const std::string& id, const Object& data, size_t index) : id(id), data(data), index(index) {}
    std::string id;
    Object data;
    size_t index;
};

struct Packet
{
    BufferChunk payload;
    std::map<std::string, Object> metadata;
    std::vector<Label> labels;
};

struct BufferManagerArgs { size_t bufferSize = 8192; size_t numBuffers = 4; };
class BufferManager
{
public:
    typedef std::shared_ptr<BufferManager> Sptr;
    static Sptr make(const std::string&, const BufferManagerArgs&) { return Sptr(); }
};

class InputPort
{
public:
    // scheduler side
    void feed(const DType& t, const void* p, size_t numElements) { _dtype = t; _base = static_cast<const char*>(p); _avail = numElements; _taken = 0; }
    void pushMessage(const Object& m) { _msgs.push_back(m); }
    size_t totalConsumed() const { return _taken; }
    // block side
    bool hasMessage() const { return !_msgs.empty(); }
    Object popMessage() { Object m = _msgs.front(); _msgs.pop_front(); return m; }
    size_t elements() const { return _avail; }
    BufferChunk buffer() const { return BufferChunk::view(_dtype, const_cast<char*>(_base) + _taken * _dtype.size(), _avail); }
    void consume(size_t n)
    {
        if (n > _avail) throw Exception("Pothos stub InputPort::consume", "more than available");
        _taken += n; _avail -= n;
    }
    void setReserve(size_t n) { reserve = n; }
    size_t reserve = 0;

private:
    DType _dtype;
    const char* _base = nullptr;
    size_t _avail = 0, _taken = 0;
    std::deque<Object> _msgs;
};

class OutputPort
{
public:
    // scheduler side
    void attach(const DType& t, void* p, size_t numElements) { _dtype = t; _base = static_cast<char*>(p); _cap = numElements; }
    std::deque<Object> messages;
    std::vector<Label> labels;   // labels posted since the scheduler last cleared them
    size_t produced = 0;         // elements produced by the last work() call (the scheduler resets it)
    // block side
    BufferChunk buffer() const { return BufferChunk::view(_dtype, _base, _cap); }
    void produce(size_t n) { produced += n; }
    template <typename T> void postMessage(const T& m) { messages.push_back(Object(m)); }
    void postLabel(const Label& l) { labels.push_back(l); }
    void setReserve(size_t n) { reserve = n; }
    size_t reserve = 0;

private:
    DType _dtype;
    char* _base = nullptr;
    size_t _cap = 0;
};

class Block
{
public:
    virtual ~Block() {}
    virtual void activate() {}
    virtual void deactivate() {}
    virtual void work() {}
    virtual BufferManager::Sptr getInputBufferManager(const std::string&, const std::string&) { return BufferManager::Sptr(); }
    virtual BufferManager::Sptr getOutputBufferManager(const std::string&, const std::string&) { return BufferManager::Sptr(); }

    template <typename... A> void registerCall(A&&...) {}
    void registerSignal(const std::string&) {}
    template <typename... A> void emitSignal(const std::string& name, const A&... a) { record(name, a...); }

    void setupInput(const std::string& name, const DType& = DType()) { _in[name]; }
    void setupInput(size_t i, const DType& t = DType()) { setupInput(std::to_string(i), t); }
    void setupOutput(const std::string& name, const DType& = DType()) { _out[name]; }
    void setupOutput(size_t i, const DType& t = DType()) { setupOutput(std::to_string(i), t); }
    InputPort* input(const std::string& name) { return &_in.at(name); }
    InputPort* input(size_t i) { return input(std::to_string(i)); }
    OutputPort* output(const std::string& name) { return &_out.at(name); }
    OutputPort* output(size_t i) { return output(std::to_string(i)); }

    // what the block told the world through emitSignal(name, value): kept as doubles, in order
    std::vector<std::pair<std::string, double>> signals;

private:
    void record(const std::string& name) { signals.push_back(std::make_pair(name, 0.0)); }
    template <typename T> void record(const std::string& name, const T& v) { signals.push_back(std::make_pair(name, double(v))); }
    std::map<std::string, InputPort> _in;
    std::map<std::string, OutputPort> _out;
};

class BlockRegistry
{
public:
    template <typename F> BlockRegistry(const std::string&, F) {}
};

} // namespace Pothos
