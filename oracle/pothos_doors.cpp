// pothos_doors.cpp -- extern "C" doors into the UNMODIFIED upstream Pothos blocks, for oracle validation only.
//
// TEST INFRASTRUCTURE ONLY.  This file contains no LoRa algorithm.  oracle/Makefile compiles it four times, each
// time with one of -DDOOR_DECODER / -DDOOR_ENCODER / -DDOOR_DEMOD / -DDOOR_MOD, and each variant #includes ONE block
// source exactly as it lies under /root/reference/LoRa-SDR (found through -I; nothing is copied into this
// repository), on top of the header-only stand-in oracle/pothos_stub/Pothos/Framework.hpp.  What is written here is
// the scheduler's side: put the input on the block's ports, call activate() once and work() until the block is done,
// collect what it posted.  The result is oracle/_ref/liblora_pothos_ref.so; tests/test_oracle_vs_pothos.py pins
// oracle/chain.c and oracle/stream.c to it bit for bit.
//
// Every block object is placement-constructed into zero-filled memory.  LoRaDemod's constructor leaves _prevValue,
// _freqError, _finefreqError and _symCount uninitialised (LoRaDemod.cpp:68-75,381-390); under a Pothos scheduler
// they hold whatever the heap held.  Zero is the one defined reading and it is what oracle/stream.c and the device
// kernel assume; DESIGN.md section 6 records it.
//
// The signatures mirror lora_oracle.h (orc_X <-> pothos_X) so the tests call both with the same arguments.
#include <Pothos/Framework.hpp>

#include <cstdlib>
#include <new>

#include "lora_oracle.h"

typedef std::complex<float> cf;

// Zero-filled, slightly over-sized heap blocks for everything this library allocates (the Makefile links it with
// -Bsymbolic, so only this library sees these replacements).  Reason: LoRaEncoder::work reads PAST THE END of its
// `bytes` vector when the last interleaver block is padded -- encodeFec is asked for numCodewords-PPM nibbles
// (LoRaEncoder.cpp:216) although fewer payload nibbles exist, up to (PPM-1)/2+1 bytes beyond bytes.size().  Under
// Pothos those padding nibbles are heap garbage (the decoder never looks at them); here they read as zero, the one
// defined value, which is what oracle/chain.c and the device encoder use.  DESIGN.md section 6 records it.
#ifdef DOOR_ENCODER // defined once for the whole library
static void* zeroed_block(std::size_t n)
{
    void* p = std::calloc(1, n + 16);
    if (!p) throw std::bad_alloc();
    return p;
}
void* operator new(std::size_t n) { return zeroed_block(n); }
void* operator new[](std::size_t n) { return zeroed_block(n); }
void operator delete(void* p) noexcept { std::free(p); }
void operator delete[](void* p) noexcept { std::free(p); }
void operator delete(void* p, std::size_t) noexcept { std::free(p); }
void operator delete[](void* p, std::size_t) noexcept { std::free(p); }
#endif

template <typename B> struct Zeroed // a block living in calloc'ed memory
{
    void* mem;
    B* b;
    template <typename... A> explicit Zeroed(A... a) : mem(std::calloc(1, sizeof(B))), b(new (mem) B(a...)) {}
    ~Zeroed() { b->~B(); std::free(mem); }
    B* operator->() { return b; }
};

static const char* cr_name(unsigned rdd)
{
    static const char* names[] = {"4/4", "4/5", "4/6", "4/7", "4/8"};
    return rdd < 5 ? names[rdd] : "?";
}

#ifdef DOOR_ENCODER
#include "LoRaEncoder.cpp"

// LoRaEncoder::work (LoRa-SDR/LoRaEncoder.cpp:161-233) on one packet of bytes.
extern "C" size_t pothos_chain_encode(const orc_chain_cfg* c, const uint8_t* payload, size_t len, uint16_t* symbols, size_t cap)
{
    try {
        Zeroed<LoRaEncoder> enc;
        enc->setSpreadFactor(c->sf);
        enc->setSymbolSize(c->ppm);
        enc->setCodingRate(cr_name(c->rdd));
        enc->enableWhitening(c->whitening != 0);
        enc->enableExplicit(c->explicit_header != 0);
        enc->enableCrc(c->crc != 0);
        enc->activate();
        Pothos::Packet pkt;
        pkt.payload = Pothos::BufferChunk(typeid(uint8_t), len);
        if (len) std::memcpy(pkt.payload.as<void*>(), payload, len);
        enc->input(0)->pushMessage(Pothos::Object(pkt));
        enc->work();
        auto& out = enc->output(0)->messages;
        if (out.empty()) return 0;
        const Pothos::Packet& r = out.front().extract<Pothos::Packet>();
        const size_t n = r.payload.elements();
        if (symbols && n <= cap) std::memcpy(symbols, r.payload.as<const void*>(), n * sizeof(uint16_t));
        return n;
    } catch (const std::exception&) {
        return size_t(-1);
    }
}
#endif

#ifdef DOOR_DECODER
#include "LoRaDecoder.cpp"

// LoRaDecoder::work (LoRa-SDR/LoRaDecoder.cpp:196-397) on one packet of symbols.  Returns the bytes posted, -1 when
// the block counted a drop, -2 when it returned without output and without a drop (:199,:208), -100 on an exception.
// Only what the block shows the outside world can be reported: `info` gets drop_line = -1/0 and nothing else.
extern "C" long pothos_chain_decode(const orc_chain_cfg* c, const uint16_t* symbols, size_t nsym, uint8_t* out, size_t cap,
                                    orc_chain_info* info)
{
    if (info) std::memset(info, 0, sizeof(*info));
    try {
        Zeroed<LoRaDecoder> dec;
        dec->setSpreadFactor(c->sf);
        dec->setSymbolSize(c->ppm);
        dec->setCodingRate(cr_name(c->rdd));
        dec->enableWhitening(c->whitening != 0);
        dec->enableCrcc(c->crc != 0);
        dec->enableInterleaving(c->interleaving != 0);
        dec->enableExplicit(c->explicit_header != 0);
        dec->enableHdr(c->hdr_out != 0);
        dec->setDataLength(c->data_length);
        dec->enableErrorCheck(c->error_check != 0);
        dec->activate();
        Pothos::Packet pkt;
        pkt.payload = Pothos::BufferChunk(typeid(uint16_t), nsym);
        if (nsym) std::memcpy(pkt.payload.as<void*>(), symbols, nsym * sizeof(uint16_t));
        dec->input(0)->pushMessage(Pothos::Object(pkt));
        dec->work();
        if (dec->getDropped() != 0) {
            if (info) info->drop_line = -1;
            return -1;
        }
        auto& posted = dec->output(0)->messages;
        if (posted.empty()) return -2;
        const Pothos::Packet& r = posted.front().extract<Pothos::Packet>();
        const size_t n = r.payload.length;
        if (n <= cap) std::memcpy(out, r.payload.as<const void*>(), n);
        return long(n);
    } catch (const std::exception&) {
        return -100;
    }
}
#endif

#ifdef DOOR_DECODER
#include <chrono>
#include <thread>
// LoRaDecoder::work over a batch, for bench.py's CPU leg: `threads` host threads, ONE block per thread configured
// once, a packet message pushed and work() called per packet.  n_bytes[p] = bytes posted, -1 dropped, -2 no output;
// out (payload_stride bytes per packet) receives the posted bytes.  Returns the seconds of the timed loops.
extern "C" double pothos_chain_decode_batch(const orc_chain_cfg* c, const uint16_t* symbols, size_t packets, size_t nsym,
                                            size_t symbol_stride, uint8_t* out, size_t payload_stride, long* n_bytes,
                                            unsigned threads)
{
    if (!c || !symbols || !n_bytes || threads == 0) return -1.0;
    if (threads > packets) threads = packets ? unsigned(packets) : 1u;
    auto body = [&](unsigned t) {
        Zeroed<LoRaDecoder> dec;
        dec->setSpreadFactor(c->sf);
        dec->setSymbolSize(c->ppm);
        dec->setCodingRate(cr_name(c->rdd));
        dec->enableWhitening(c->whitening != 0);
        dec->enableCrcc(c->crc != 0);
        dec->enableInterleaving(c->interleaving != 0);
        dec->enableExplicit(c->explicit_header != 0);
        dec->enableHdr(c->hdr_out != 0);
        dec->setDataLength(c->data_length);
        dec->enableErrorCheck(c->error_check != 0);
        dec->activate();
        unsigned long long dropped = 0;
        for (size_t p = packets * t / threads; p < packets * (t + 1) / threads; ++p) {
            auto& posted = dec->output(0)->messages;
            try {
                Pothos::Packet pkt;
                pkt.payload = Pothos::BufferChunk(typeid(uint16_t), nsym);
                std::memcpy(pkt.payload.as<void*>(), symbols + p * symbol_stride, nsym * sizeof(uint16_t));
                dec->input(0)->pushMessage(Pothos::Object(pkt));
                dec->work();
                if (dec->getDropped() != dropped) { dropped = dec->getDropped(); n_bytes[p] = -1; }
                else if (posted.empty()) n_bytes[p] = -2;
                else {
                    const Pothos::Packet& r = posted.front().extract<Pothos::Packet>();
                    n_bytes[p] = long(r.payload.length);
                    if (out) std::memcpy(out + p * payload_stride, r.payload.as<const void*>(),
                                         r.payload.length < payload_stride ? r.payload.length : payload_stride);
                }
            } catch (const std::exception&) {
                // a noise packet whose header announces no CRC and fewer than two bytes makes the block ask for a
                // (size_t)-wrapped buffer (LoRaDecoder.cpp:381-392): undefined in the reference, an exception here
                n_bytes[p] = -100;
            }
            dec->signals.clear();
            posted.clear();
        }
    };
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < threads; ++t) pool.emplace_back(body, t);
    body(0);
    for (auto& th : pool) th.join();
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}
#endif

#ifdef DOOR_MOD
#include "LoRaMod.cpp"

// LoRaMod::work (LoRa-SDR/LoRaMod.cpp:109-238) for one packet of symbols: work() is called until the block posts its
// "txEnd" label (:226).  Returns the samples produced (out may be NULL to size the buffer).
extern "C" size_t pothos_frame_modulate(unsigned sf, unsigned ovs, const uint16_t* symbols, size_t n_sym, unsigned char sync,
                                        unsigned padding, float ampl, orc_c32* out)
{
    try {
        const size_t sf_arg = sf;
        Zeroed<LoRaMod> mod(sf_arg);
        mod->setSync(sync);
        mod->setPadding(padding);
        mod->setAmplitude(ampl);
        mod->setOvs(ovs);
        mod->activate();
        const size_t NN = (size_t(1) << sf) * ovs;
        std::vector<cf> slab(NN);
        Pothos::OutputPort* port = mod->output(0);
        port->attach(typeid(cf), slab.data(), NN);
        Pothos::Packet pkt;
        pkt.payload = Pothos::BufferChunk(typeid(uint16_t), n_sym);
        if (n_sym) std::memcpy(pkt.payload.as<void*>(), symbols, n_sym * sizeof(uint16_t));
        mod->input(0)->pushMessage(Pothos::Object(pkt));
        size_t pos = 0;
        for (size_t calls = 0; calls < n_sym + padding + 64; ++calls) {
            port->produced = 0;
            port->labels.clear();
            mod->work();
            if (out) std::memcpy(reinterpret_cast<cf*>(out) + pos, slab.data(), port->produced * sizeof(cf));
            pos += port->produced;
            bool end = false;
            for (const auto& l : port->labels) end = end || l.id == "txEnd";
            if (end) break;
        }
        return pos;
    } catch (const std::exception&) {
        return size_t(-1);
    }
}
#endif

#ifdef DOOR_DEMOD
#include "LoRaDemod.cpp"

// LoRaDemod::work (LoRa-SDR/LoRaDemod.cpp:145-327) over one finite buffer: the scheduler calls work() while at
// least 2N samples are buffered (:148) and the block consumes what it decides (:320).  Packets come out of port 0;
// pkt_start is the input position after the call that posted the "QC" label (:283, the first data window), pkt_err
// the value of the last "error" signal (:267).  Same outputs as orc_stream_demod.
extern "C" size_t pothos_stream_demod(unsigned sf, unsigned char sync, float thresh, size_t mtu, const orc_c32* samples,
                                      size_t count, int16_t* out_symbols, size_t sym_cap, uint32_t* pkt_len,
                                      uint64_t* pkt_start, int32_t* pkt_err, size_t pkt_cap, uint64_t* consumed)
{
    try {
        const size_t sf_arg = sf;
        Zeroed<LoRaDemod> dem(sf_arg);
        dem->setSync(sync);
        dem->setThreshold(thresh);
        dem->setMTU(mtu);
        const size_t N = size_t(1) << sf;
        std::vector<cf> raw(2 * N), dec(2 * N), fft(N);
        dem->output("raw")->attach(typeid(cf), raw.data(), 2 * N);
        dem->output("dec")->attach(typeid(cf), dec.data(), 2 * N);
        dem->output("fft")->attach(typeid(cf), fft.data(), N);
        Pothos::InputPort* in = dem->input(0);
        Pothos::OutputPort* out0 = dem->output(0);
        Pothos::OutputPort* rawp = dem->output("raw");
        in->feed(typeid(cf), samples, count);
        dem->activate();
        size_t n_pkt = 0, sym_used = 0;
        uint64_t cur_start = 0;
        while (in->elements() >= 2 * N) {
            rawp->labels.clear();
            dem->work();
            for (const auto& l : rawp->labels)
                if (l.id == "QC") cur_start = in->totalConsumed();
            while (!out0->messages.empty()) {
                const Pothos::Packet& p = out0->messages.front().extract<Pothos::Packet>();
                const size_t n = p.payload.length / sizeof(int16_t);
                const int16_t* s = p.payload.as<const int16_t*>();
                for (size_t k = 0; k < n && sym_used + k < sym_cap; ++k) out_symbols[sym_used + k] = s[k];
                if (n_pkt < pkt_cap) {
                    pkt_len[n_pkt] = uint32_t(n);
                    pkt_start[n_pkt] = cur_start;
                    double err = 0;
                    for (const auto& sg : dem->signals) if (sg.first == "error") err = sg.second;
                    pkt_err[n_pkt] = int32_t(err);
                }
                ++n_pkt;
                sym_used += n;
                if (sym_used > sym_cap) sym_used = sym_cap;
                out0->messages.pop_front();
            }
            dem->signals.erase(dem->signals.begin(), dem->signals.size() > 16 ? dem->signals.end() - 16 : dem->signals.begin());
        }
        if (consumed) *consumed = in->totalConsumed();
        return n_pkt;
    } catch (const std::exception&) {
        return size_t(-1);
    }
}
#endif
