// m17-demod -- the reference CLI's surface (apps/m17-demod.cpp:372-499) on the GPU path.
//
//   m17-demod [-i] [-b] [-l] [-v] [-d] [-q] [-h] [-V] [--frames] < baseband.s16
//
// stdin: int16 LE mono 48 kS/s.  stderr: the reference's LSF line (:131-152), EOS /
// LICH notices, packet messages (:207-253), diagnostics line (:343-350) and BER line
// (:357-368).  stdout: decoded frame payloads -- per stream frame the 16 Codec2 bytes
// (Codec2 voice synthesis is outside this project), reassembled packets for packet
// mode; with --frames one text record per frame callback in the format of
// oracle/_ref/m17-demod-ref, which is what the parity tests diff.
#include "m17cxx_b200/M17Demodulator.h"
#include "m17cxx_b200/ax25_frame.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iomanip>
#include <iostream>
#include <string>
#include <vector>

namespace {

const char VERSION[] = "2.2";

bool display_lsf = false, invert_input = false, quiet = false, debug = false, noise_blanker = false, frames_mode = false;
std::vector<uint8_t> current_packet;
size_t packet_frame_counter = 0;
uint64_t n_callbacks = 0;

// PRBS9 BER accounting, Util.h:320-413
struct PRBS9 {
    uint16_t state = 1; bool synced = false; uint8_t sync_count = 0; uint32_t bit_count = 0, err_count = 0;
    uint8_t history[16] = {0}; size_t hist_count = 0, hist_pos = 0;
    bool generate() { bool r = ((state >> 8) ^ (state >> 4)) & 1; state = uint16_t(((state << 1) | r) & 0x1FF); return r; }
    void count_errors(bool error) {
        bit_count += 1;
        hist_count -= (history[hist_pos >> 3] & (1 << (hist_pos & 7))) != 0;
        if (error) { err_count += 1; hist_count += 1; history[hist_pos >> 3] |= uint8_t(1 << (hist_pos & 7)); if (hist_count >= 25) synced = false; }
        else history[hist_pos >> 3] &= uint8_t(~(1 << (hist_pos & 7)));
        if (++hist_pos == 128) hist_pos = 0;
    }
    bool synchronize(bool bit) {
        bool r = (bit ^ (state >> 8) ^ (state >> 4)) & 1;
        state = uint16_t(((state << 1) | bit) & 0x1FF);
        if (r) sync_count = 0;
        else if (++sync_count == 18) { synced = true; bit_count += 18; std::memset(history, 0, 16); hist_count = hist_pos = 0; sync_count = 0; }
        return r;
    }
    bool validate(bool bit) { if (!synced) return synchronize(bit); bool r = bit ^ generate(); count_errors(r); return r; }
    void reset() { *this = PRBS9(); }
} prbs;

// CRC-16/X-25 residue check used for packets, apps/m17-demod.cpp:219-223 (0x0f47)
uint16_t crc_x25(const uint8_t* d, size_t n)
{
    uint16_t reg = 0xFFFF;
    for (size_t i = 0; i != n; ++i) { reg ^= d[i]; for (int b = 0; b != 8; ++b) reg = (reg & 1) ? uint16_t((reg >> 1) ^ 0x8408) : uint16_t(reg >> 1); }
    return uint16_t(~reg);
}

// LinkSetupFrame.h:95-121
std::string decode_callsign(const uint8_t* enc)
{
    static const char map[] = "xABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789-/.";
    bool bcast = true;
    for (int i = 0; i != 6; ++i) bcast = bcast && enc[i] == 0xFF;
    if (bcast) return "BROADCAST";
    uint64_t v = 0;
    for (int i = 0; i != 6; ++i) v = (v << 8) | enc[i];
    std::string r;
    while (v) { r.push_back(map[v % 40]); v /= 40; }
    return r;
}

// apps/m17-demod.cpp:80-175
void dump_lsf(const std::array<uint8_t, 30>& lsf)
{
    if (display_lsf) {
        std::cerr << "\nSRC: " << decode_callsign(&lsf[6]) << ", DEST: " << decode_callsign(&lsf[0]);
        uint16_t type = uint16_t((lsf[12] << 8) | lsf[13]);
        static const char* str[] = {"UNK", "D/D", "V/V", "V/D"};
        static const char* pkt[] = {"UNK", "RAW", "ENC", "UNK"};
        std::cerr << ", " << ((type & 1) ? "STR:" : "PKT:") << ((type & 1) ? str[(type & 6) >> 1] : pkt[(type & 6) >> 1]);
        std::cerr << " CAN:" << std::dec << std::setw(2) << std::setfill('0') << int((type & 0x780) >> 7);
        std::cerr << ", NONCE: ";
        for (size_t i = 14; i != 28; ++i) std::cerr << std::hex << std::setw(2) << std::setfill('0') << int(lsf[i]);
        std::cerr << ", CRC: " << std::hex << std::setw(4) << std::setfill('0') << ((lsf[28] << 8) | lsf[29]);
        std::cerr << std::dec << std::endl;
    }
    current_packet.clear();
    packet_frame_counter = 0;
    // The reference indexes lsf[109..111] of a 30-byte array here (out of bounds); the
    // intended test is the LSF type field: bit 0 clear = packet, bits 2..1 = packet type.
    if (!(lsf[13] & 1)) {
        int packet_type = (lsf[13] >> 1) & 3;
        if (packet_type != 1) {
            if (packet_type != 2) std::cerr << "LSF for reserved packet type" << std::endl;
            current_packet.insert(current_packet.end(), lsf.begin(), lsf.end());
        }
    }
}

// apps/m17-demod.cpp:207-253
bool decode_packet(const std::array<uint8_t, 26>& seg)
{
    if (seg[25] & 0x80) {
        size_t packet_size = std::min<size_t>((seg[25] & 0x7F) >> 2, 25);
        current_packet.insert(current_packet.end(), seg.begin(), seg.begin() + packet_size);
        uint16_t checksum = crc_x25(current_packet.data(), current_packet.size());
        // boost::crc_optimal<16,0x1021,0xFFFF,0xFFFF,true,true> over data+FCS leaves 0x0f47
        if (checksum == 0x0f47) {
            // the reference prints the packet as an AX.25 frame on stderr (:224-229); the packet itself also goes to
            // stdout here (this CLI's stdout carries the decoded payloads)
            mobilinkd::ax25_frame frame(std::string(current_packet.begin(), current_packet.end()));
            std::cerr << '\n';
            mobilinkd::write(std::cerr, frame);
            std::cout.write(reinterpret_cast<const char*>(current_packet.data()), std::streamsize(current_packet.size()));
            return true;
        }
        std::cerr << "\nPacket checksum error: " << std::hex << checksum << std::dec << std::endl;
        return false;
    }
    size_t frame_number = (seg[25] & 0x7F) >> 2;
    if (frame_number != packet_frame_counter) {
        std::cerr << "\nPacket frame sequence error. Got " << frame_number << ", expected " << packet_frame_counter << "\n";
        return false;
    }
    packet_frame_counter += 1;
    current_packet.insert(current_packet.end(), seg.begin(), seg.begin() + 25);
    return true;
}

void print_record(const mobilinkd::M17FrameDecoder::output_buffer_t& f, int cost)
{
    static const char* names[] = {"LSF", "LICH", "STREAM", "BASIC_PACKET", "FULL_PACKET", "BERT"};
    static const int plen[] = {0, 6, 18, 26, 26, 25};
    int t = int(f.type);
    std::printf("0 %llu - %s cost=%d payload=", (unsigned long long) n_callbacks, names[t], cost);
    for (int i = 0; i != plen[t]; ++i) std::printf("%02x", f.packet[size_t(i)]);
    std::printf(" lsf=");
    for (int i = 0; i != 30; ++i) std::printf("%02x", f.lsf[size_t(i)]);
    std::printf("\n");
}

// apps/m17-demod.cpp:307-336
bool handle_frame(const mobilinkd::M17FrameDecoder::output_buffer_t& frame, int viterbi_cost)
{
    using FrameType = mobilinkd::M17FrameDecoder::FrameType;
    if (frames_mode) { print_record(frame, viterbi_cost); ++n_callbacks; return true; }
    ++n_callbacks;
    switch (frame.type) {
    case FrameType::LSF: dump_lsf(frame.lsf); return true;
    case FrameType::LICH: std::cerr << "\nLICH" << std::endl; return true;
    case FrameType::STREAM: {
        bool result = true;
        if (viterbi_cost < 70 && (frame.stream[0] & 0x80)) { if (display_lsf) std::cerr << "\nEOS" << std::endl; result = false; }
        std::array<uint8_t, 16> codec{};
        if (!(noise_blanker && viterbi_cost > 80)) std::copy(frame.stream.begin() + 2, frame.stream.end(), codec.begin());
        std::cout.write(reinterpret_cast<const char*>(codec.data()), 16);     // 2 x Codec2 3200 frames
        return result;
    }
    case FrameType::BASIC_PACKET:
    case FrameType::FULL_PACKET: return decode_packet(frame.packet);
    case FrameType::BERT:
        for (int j = 0; j != 24; ++j) { uint8_t b = frame.bert[size_t(j)]; for (int i = 0; i != 8; ++i) { prbs.validate(b & 0x80); b = uint8_t(b << 1); } }
        { uint8_t b = frame.bert[24]; for (int i = 0; i != 5; ++i) { prbs.validate(b & 0x80); b = uint8_t(b << 1); } }
        return true;
    }
    return true;
}

// apps/m17-demod.cpp:338-370
void diagnostic_callback(bool dcd, float evm, float deviation, float offset, bool locked, float clock, int sample_index,
                         int sync_index, int clock_index, int viterbi_cost)
{
    if (debug) {
        std::cerr << "\rdcd: " << std::setw(1) << int(dcd)
                  << ", evm: " << std::setfill(' ') << std::setprecision(2) << std::setw(6) << evm * 100 << "%"
                  << ", deviation: " << std::setw(5) << int(deviation)
                  << "Hz, freq offset: " << std::setfill(' ') << std::setw(5) << int(offset * 800)
                  << "Hz, locked: " << std::boolalpha << std::setw(5) << locked << std::dec
                  << ", clock: " << std::setprecision(2) << std::fixed << std::setw(8) << (clock * 1'000'000.0)
                  << "ppm, sample: " << std::setw(1) << sample_index << ", " << sync_index << ", " << clock_index
                  << ", cost: " << std::setfill(' ') << std::setw(3) << viterbi_cost;
    }
    // (prbs.bits() of the reference asserts `synced`; built with NDEBUG, as its CMake release build is, it is bit_count)
    if (!dcd && prbs.bit_count > 0) prbs.reset();
    if (prbs.bit_count > 0 && !quiet) {
        std::cerr << (debug ? ", " : "\r");
        char buffer[40];
        std::snprintf(buffer, 40, "BER: %-1.6lf (%lu bits)", double(prbs.err_count) / double(prbs.bit_count), (unsigned long) prbs.bit_count);
        std::cerr << buffer;
    }
    std::cerr << std::flush;
}

void usage()
{
    std::cout << "Read M17 baseband from STDIN and write decoded frame payloads to STDOUT\n"
                 "Program options:\n"
                 "  -h [ --help ]           Print this help message and exit.\n"
                 "  -V [ --version ]        Print the application verion and exit.\n"
                 "  -i [ --invert ]         invert the received baseband\n"
                 "  -b [ --noise-blanker ]  noise blanker -- silence likely corrupt audio\n"
                 "  -l [ --lsf ]            display the decoded LSF\n"
                 "  -v [ --verbose ]        verbose output\n"
                 "  -d [ --debug ]          debug-level output\n"
                 "  -q [ --quiet ]          silence all output -- no BERT output\n"
                 "  --frames                one text record per decoded frame on STDOUT\n"
                 "  --fma                   fused multiply-add matched filter (not bit-identical)\n"
              << std::endl;
}

} // namespace

int main(int argc, char* argv[])
{
    bool verbose = false, fma = false;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "-h" || a == "--help") { usage(); return 0; }
        else if (a == "-V" || a == "--version") { std::cout << argv[0] << ": " << VERSION << std::endl; return 0; }
        else if (a == "-i" || a == "--invert") invert_input = true;
        else if (a == "-b" || a == "--noise-blanker") noise_blanker = true;
        else if (a == "-l" || a == "--lsf") display_lsf = true;
        else if (a == "-v" || a == "--verbose") verbose = true;
        else if (a == "-d" || a == "--debug") debug = true;
        else if (a == "-q" || a == "--quiet") quiet = true;
        else if (a == "--frames") frames_mode = true;
        else if (a == "--fma") fma = true;
        else if (a == "--ax25-print" && i + 1 < argc) {
            // test hook: print the file's bytes the way a reassembled packet is printed (no device needed)
            std::FILE* f = std::fopen(argv[++i], "rb");
            if (!f) return EXIT_FAILURE;
            std::string bytes;
            int ch;
            while ((ch = std::fgetc(f)) != EOF) bytes.push_back(char(ch));
            std::fclose(f);
            mobilinkd::write(std::cout, mobilinkd::ax25_frame(bytes));
            return EXIT_SUCCESS;
        }
        else { std::cerr << "unrecognised option '" << a << "'" << std::endl; usage(); return 0; }
    }
    if (int(debug) + int(verbose) + int(quiet) > 1) {
        std::cerr << "Only one of quiet, verbos or debug may be chosen." << std::endl;
        return 0;
    }

    try {
        uint32_t flags = (invert_input ? M17B_INVERT : 0u) | (fma ? M17B_FIR_FMA : 0u);
        mobilinkd::M17Demodulator<float> demod(handle_frame, 0, 48000, flags);
        demod.diagnostics(diagnostic_callback);

        // apps/m17-demod.cpp:484-490: `while (std::cin)` tests the stream before the read,
        // so the last sample is processed a second time at EOF.
        std::vector<int16_t> buf(1 << 16);
        int16_t last = 0;
        bool any = false;
        size_t got;
        while ((got = std::fread(buf.data(), 2, buf.size(), stdin)) > 0) {
            for (size_t i = 0; i != got; ++i) demod.push(buf[i]);
            last = buf[got - 1];
            any = true;
        }
        demod.push(any ? last : int16_t(0));
        demod.flush();
    } catch (const std::exception& ex) {
        std::cerr << "m17-demod: " << ex.what() << std::endl;
        return EXIT_FAILURE;
    }
    std::cerr << std::endl;
    return EXIT_SUCCESS;
}
