// ax25_frame.h -- AX.25 UI-frame pretty printer for reassembled M17 packets.
//
// Replaces include/m17cxx/ax25_frame.h:22-263 (class mobilinkd::ax25_frame and write(std::ostream&, ...)),
// which apps/m17-demod.cpp:224-229 calls on every packet whose CRC checks.  Same class name, accessors and
// output text; written from the reference's observable behaviour:
//   * addresses are 7 bytes: six characters shifted left by one, then the SSID byte; an address is printed up to
//     its first space (six characters at most) followed by "-<ssid>" when the SSID (low 4 bits of the shifted
//     seventh byte) is not zero;
//   * bit 0 of an address's last byte clear = another address follows (the source's decides whether there are
//     repeaters, each repeater's whether there is one more -- as long as seven more bytes are there);
//   * the control byte's low two bits give the type (0, 2: information, 1: supervisory, 3: unnumbered); only
//     unnumbered frames carry a PID byte; the information field runs to the two FCS bytes;
//   * a frame shorter than 17 bytes parses to nothing; one that ends before control + PID + FCS keeps only its
//     addresses.
#pragma once

#include <cstdint>
#include <iomanip>
#include <optional>
#include <ostream>
#include <string>
#include <vector>

namespace mobilinkd {

struct ax25_frame
{
    using repeaters_type = std::vector<std::string>;
    using pid_type = std::optional<uint8_t>;
    enum frame_type { UNDEFINED, INFORMATION, SUPERVISORY, UNNUMBERED };

    explicit ax25_frame(const std::string& frame) { parse(frame); }

    std::string destination() const { return destination_; }
    std::string source() const { return source_; }
    repeaters_type repeaters() const { return repeaters_; }
    frame_type type() const { return type_; }
    std::string info() const { return info_; }
    uint16_t fcs() const { return fcs_; }
    uint16_t crc() const { return 0; }
    pid_type pid() const { return pid_; }

private:
    std::string destination_, source_, info_;
    repeaters_type repeaters_;
    frame_type type_ = UNDEFINED;
    uint16_t fcs_ = 0xFFFF;
    pid_type pid_;

    // Text form of the 7-byte address at `at`; *more = another address follows it.
    static std::string address(const std::string& f, size_t at, bool* more)
    {
        *more = (uint8_t(f[at + 6]) & 1) == 0;
        std::string call;
        for (size_t i = 0; i != 6; ++i) {
            const char c = char(uint8_t(f[at + i]) >> 1);
            if (c == ' ') break;
            call.push_back(c);
        }
        const int ssid = (uint8_t(f[at + 6]) >> 1) & 0x0F;
        if (ssid) call += "-" + std::to_string(ssid);
        return call;
    }

    void parse(const std::string& f)
    {
        if (f.size() < 17) return;
        // the FCS as transmitted: low byte first, every bit mirrored
        const uint16_t raw = uint16_t((uint8_t(f[f.size() - 1]) << 8) | uint8_t(f[f.size() - 2]));
        uint16_t mirrored = 0;
        for (int b = 0; b != 16; ++b) mirrored = uint16_t((mirrored << 1) | ((raw >> b) & 1));
        fcs_ = mirrored;

        bool more = false;
        destination_ = address(f, 0, &more);
        source_ = address(f, 7, &more);
        size_t at = 14;
        while (more && at + 7 < f.size()) {
            repeaters_.push_back(address(f, at, &more));
            at += 7;
        }
        if (f.size() < at + 5) return;
        const uint8_t control = uint8_t(f[at++]);
        switch (control & 3) {
        case 1: type_ = SUPERVISORY; break;
        case 3: type_ = UNNUMBERED; break;
        default: type_ = INFORMATION; break;
        }
        if (type_ == UNNUMBERED) pid_ = uint8_t(f[at++]);
        info_.assign(f.begin() + std::string::difference_type(at), f.end() - 2);
    }
};

inline void write(std::ostream& os, const ax25_frame& frame)
{
    os << "Dest: " << frame.destination() << std::endl << "Source: " << frame.source() << std::endl;
    const auto via = frame.repeaters();
    if (!via.empty()) {
        os << "Via: ";
        for (const auto& r : via) os << r << " ";
        os << std::endl;
    }
    if (frame.pid()) os << "PID: " << std::setbase(16) << int(*frame.pid()) << std::endl;     // (the base sticks, like the reference's)
    os << "Info: " << std::endl << frame.info() << std::endl;
}

} // namespace mobilinkd
