// h5lite_impl.hpp — included by h5lite.hpp.  Section numbers refer to the HDF5 File Format
// Specification version 3.0.
#pragma once

namespace h5lite {

class File {
  public:
    explicit File(const std::string& path) : path_(path)
    {
        fd_ = ::open(path.c_str(), O_RDONLY | O_CLOEXEC);
        if (fd_ < 0) throw Error("cannot open " + path);
        struct stat st;
        if (::fstat(fd_, &st) != 0 || st.st_size < 8) {
            ::close(fd_);
            throw Error(path + ": not an HDF5 file (too short)");
        }
        size_ = (uint64_t)st.st_size;
        void* m = ::mmap(nullptr, (size_t)size_, PROT_READ, MAP_PRIVATE, fd_, 0);
        if (m == MAP_FAILED) {
            ::close(fd_);
            throw Error("cannot map " + path);
        }
        map_ = (const uint8_t*)m;
        try {
            superblock();
        } catch (...) {
            ::munmap((void*)map_, (size_t)size_);
            ::close(fd_);
            throw;
        }
    }
    ~File()
    {
        if (map_) ::munmap((void*)map_, (size_t)size_);
        if (fd_ >= 0) ::close(fd_);
    }
    File(const File&) = delete;
    File& operator=(const File&) = delete;

    // object header address of the dataset or group at `path` ("/a/b" or "a/b")
    uint64_t find(const std::string& path) const
    {
        uint64_t obj = root_;
        size_t i = 0;
        while (i < path.size()) {
            while (i < path.size() && path[i] == '/') ++i;
            size_t j = i;
            while (j < path.size() && path[j] != '/') ++j;
            if (j > i) {
                const std::string name = path.substr(i, j - i);
                obj = lookup(obj, name);
                if (obj == kUndef) throw Error("Unable to open the dataset \"" + path + "\": no link named \"" + name + "\" in " + path_);
            }
            i = j;
        }
        return obj;
    }

    bool exists(const std::string& path) const
    {
        try {
            find(path);
            return true;
        } catch (const Error&) {
            return false;
        }
    }

    Dataset dataset(const std::string& path) const { return describe(find(path), path); }

    template <class T>
    void read(const std::string& path, std::vector<T>& out) const
    {
        const Dataset d = dataset(path);
        std::vector<uint8_t> raw;
        read_raw(d, raw, path);
        convert(d.type, raw.data(), d.count(), out, path);
    }

    template <class T>
    void read_scalar(const std::string& path, T& out) const
    {
        std::vector<T> v;
        read(path, v);
        if (v.size() != 1) throw Error("\"" + path + "\" holds " + std::to_string(v.size()) + " elements where a scalar is expected");
        out = v[0];
    }

    int superblock_version() const { return sb_version_; }

  private:
    Cur at(uint64_t off) const
    {
        if (off == kUndef) throw Error("undefined address followed");
        return Cur(map_, size_, base_ + off);
    }
    uint64_t rd_off(Cur& c) const
    {
        uint64_t v = c.u(so_);
        if (so_ < 8 && v == ((1ull << (8 * so_)) - 1)) v = kUndef;
        return v;
    }
    uint64_t rd_len(Cur& c) const { return c.u(sl_); }

    // II.A: the signature sits at 0 or at 512 * 2^k
    void superblock()
    {
        static const uint8_t sig[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
        uint64_t off = 0;
        for (;;) {
            if (off + 8 > size_) throw Error(path_ + ": not an HDF5 file (no signature)");
            if (std::memcmp(map_ + off, sig, 8) == 0) break;
            off = off == 0 ? 512 : off * 2;
        }
        Cur c(map_, size_, off + 8);
        sb_version_ = c.u8();
        if (sb_version_ == 0 || sb_version_ == 1) {
            c.skip(4);                 // free-space, root entry, reserved, shared-header versions
            so_ = c.u8();
            sl_ = c.u8();
            c.skip(1);
            c.skip(4);                 // group leaf / internal node K
            c.skip(4);                 // consistency flags
            if (sb_version_ == 1) c.skip(4);
            check_sizes();
            base_ = 0;
            const uint64_t base = rd_off(c);
            rd_off(c);                 // free-space info
            rd_off(c);                 // end of file
            rd_off(c);                 // driver information
            base_ = (base == kUndef ? 0 : base) + 0;
            // root group symbol table entry (III.C)
            rd_off(c);                 // link name offset
            root_ = rd_off(c);
        } else if (sb_version_ == 2 || sb_version_ == 3) {
            so_ = c.u8();
            sl_ = c.u8();
            c.skip(1);                 // consistency flags
            check_sizes();
            const uint64_t base = rd_off(c);
            rd_off(c);                 // superblock extension
            rd_off(c);                 // end of file
            root_ = rd_off(c);
            base_ = base == kUndef ? 0 : base;
        } else {
            throw Error("superblock version " + std::to_string(sb_version_) + " is not supported");
        }
        if (base_ == 0) base_ = 0;
        // a file with a user block has its addresses relative to the base address, which HDF5
        // sets to the superblock's own offset
        if (off != 0 && base_ == 0) base_ = off;
        if (root_ == kUndef) throw Error("the root group has no object header");
    }
    void check_sizes() const
    {
        if ((so_ != 2 && so_ != 4 && so_ != 8) || (sl_ != 2 && sl_ != 4 && sl_ != 8))
            throw Error("unsupported size of offsets / lengths");
    }

    // IV.A.1: all messages of an object header, continuation blocks followed
    std::vector<Msg> messages(uint64_t addr) const
    {
        std::vector<Msg> out;
        Cur c = at(addr);
        c.need(4);
        if (std::memcmp(map_ + c.pos, "OHDR", 4) == 0) {
            c.skip(4);
            if (c.u8() != 2) throw Error("object header version is not 2");
            const unsigned flags = c.u8();
            if (flags & 0x20u) c.skip(16);
            if (flags & 0x10u) c.skip(4);
            const uint64_t chunk0 = c.u(1 << (flags & 3u));
            std::vector<std::pair<uint64_t, uint64_t>> blocks;   // (file offset, length of message area)
            blocks.push_back({c.pos, chunk0});
            for (size_t b = 0; b < blocks.size(); ++b) {
                Cur m(map_, size_, blocks[b].first);
                const uint64_t end = blocks[b].first + blocks[b].second;
                // a message header is 4 bytes (+2 with creation order); a smaller tail is a gap
                const uint64_t hdr = 4 + ((flags & 0x04u) ? 2 : 0);
                while (m.pos + hdr <= end) {
                    const int type = m.u8();
                    const uint64_t sz = m.u16();
                    const unsigned mflags = m.u8();
                    if (flags & 0x04u) m.skip(2);
                    if (m.pos + sz > end) throw Error("object header message runs past its chunk");
                    if (type == 0x10) {
                        Cur k(map_, size_, m.pos);
                        const uint64_t coff = rd_off(k), clen = rd_len(k);
                        Cur cc = at(coff);
                        if (!cc.sig("OCHK")) throw Error("bad object header continuation signature");
                        if (clen < 8) throw Error("bad object header continuation length");
                        blocks.push_back({cc.pos, clen - 8});   // minus signature and checksum
                    } else if (type != 0) {
                        out.push_back(Msg{type, mflags, m.pos, sz});
                    }
                    m.skip(sz);
                }
            }
            return out;
        }
        // version 1 (IV.A.1.a)
        if (c.u8() != 1) throw Error("object header is neither version 1 nor version 2");
        c.skip(1);
        unsigned nmsg = c.u16();
        c.skip(4);                       // reference count
        const uint64_t hsize = c.u32();
        c.skip(4);                       // pad to 8
        std::vector<std::pair<uint64_t, uint64_t>> blocks;
        blocks.push_back({c.pos, hsize});
        for (size_t b = 0; b < blocks.size() && nmsg > 0; ++b) {
            Cur m(map_, size_, blocks[b].first);
            const uint64_t end = blocks[b].first + blocks[b].second;
            while (nmsg > 0 && m.pos + 8 <= end) {
                const int type = m.u16();
                const uint64_t sz = m.u16();
                const unsigned mflags = m.u8();
                m.skip(3);
                if (m.pos + sz > end) throw Error("object header message runs past its block");
                --nmsg;
                if (type == 0x10) {
                    Cur k(map_, size_, m.pos);
                    const uint64_t coff = rd_off(k), clen = rd_len(k);
                    blocks.push_back({base_ + coff, clen});
                } else if (type != 0) {
                    out.push_back(Msg{type, mflags, m.pos, sz});
                }
                m.skip(sz);
            }
        }
        return out;
    }

#include "h5lite_groups.inc"
#include "h5lite_data.inc"

    std::string path_;
    int fd_ = -1;
    const uint8_t* map_ = nullptr;
    uint64_t size_ = 0;
    uint64_t base_ = 0;
    uint64_t root_ = kUndef;
    int so_ = 8, sl_ = 8;
    int sb_version_ = 0;
};

}  // namespace h5lite
