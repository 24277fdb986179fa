// ImageIO.hpp — image file decoding / encoding for the host API (header only, no dependencies).
//
// The reference decodes texture files with stb_image (Texture.cpp:4-36: stbi_load for Image, stbi_loadf for FloatImage) and
// writes its film with stb_image_write (Film.hpp:172-217).  Those are third-party sources inside the reference tree; this
// file is an independent implementation of the formats this renderer needs, written from the format specifications:
//   read : PNG (8 / 16 bit, grey / grey+alpha / RGB / RGBA / palette, non-interlaced; own inflate), binary PPM / PGM,
//          BMP (uncompressed 24 / 32 bit), TGA (true colour / grey, raw or RLE), Radiance .hdr (RGBE, flat or RLE), PFM.
//          Baseline / progressive JPEG decoding is not implemented: load_u8 says so (convert the file to PNG).
//   write: baseline JPEG (JFIF, 4:4:4, standard Huffman tables, quality 1..100).
// Layout follows stb_image's default: row 0 = top row, channels interleaved, `channels` as stored in the file.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace ptb_io {

struct ImageU8 { int width = 0, height = 0, channels = 0; std::vector<uint8_t> data; };
struct ImageF32 { int width = 0, height = 0, channels = 0; std::vector<float> data; };

inline std::vector<uint8_t> read_file(const std::string& path){
    std::ifstream in(path, std::ios::binary);
    if(!in)throw std::runtime_error("cannot open image file: " + path);
    return std::vector<uint8_t>((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
}

// ---- inflate (RFC 1950 / 1951) ---------------------------------------------------------------------------------------
namespace detail {
struct BitReader {
    const uint8_t* p; size_t n, pos = 0; uint32_t acc = 0; int cnt = 0;
    uint32_t bits(int k){
        while(cnt < k){ if(pos >= n)throw std::runtime_error("inflate: out of data"); acc |= (uint32_t)p[pos++] << cnt; cnt += 8; }
        const uint32_t v = acc & ((k == 32) ? 0xFFFFFFFFu : ((1u << k) - 1u));
        acc >>= k; cnt -= k;
        return v;
    }
    void align(){ acc = 0; cnt = 0; }
};
struct Huffman {                    // canonical code: counts per length + symbols in code order
    uint16_t count[16] = {}, symbol[320] = {};
    void build(const uint8_t* len, int n){
        std::memset(count, 0, sizeof count);
        for(int i = 0; i < n; i++)count[len[i]]++;
        count[0] = 0;
        uint16_t offs[16]; offs[1] = 0;
        for(int l = 1; l < 15; l++)offs[l + 1] = (uint16_t)(offs[l] + count[l]);
        for(int i = 0; i < n; i++)if(len[i])symbol[offs[len[i]]++] = (uint16_t)i;
    }
    int decode(BitReader& br) const{
        int code = 0, first = 0, index = 0;
        for(int l = 1; l <= 15; l++){
            code |= (int)br.bits(1);
            const int c = count[l];
            if(code - c < first)return symbol[index + (code - first)];
            index += c; first += c; first <<= 1; code <<= 1;
        }
        throw std::runtime_error("inflate: bad Huffman code");
    }
};
}

inline std::vector<uint8_t> inflate_zlib(const uint8_t* src, size_t n){
    using namespace detail;
    if(n < 6 || (src[0] & 0x0F) != 8 || ((src[0] << 8 | src[1]) % 31) != 0)throw std::runtime_error("inflate: not a zlib stream");
    BitReader br { src + 2, n - 2 };
    std::vector<uint8_t> out;
    static const uint16_t lbase[29] = { 3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258 };
    static const uint8_t lext[29] = { 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0 };
    static const uint16_t dbase[30] = { 1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577 };
    static const uint8_t dext[30] = { 0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13 };
    for(bool last = false; !last; ){
        last = br.bits(1) != 0;
        const uint32_t type = br.bits(2);
        if(type == 0){
            br.align();
            if(br.pos + 4 > br.n)throw std::runtime_error("inflate: truncated stored block");
            const uint32_t len = br.p[br.pos] | br.p[br.pos + 1] << 8;
            br.pos += 4;
            if(br.pos + len > br.n)throw std::runtime_error("inflate: truncated stored block");
            out.insert(out.end(), br.p + br.pos, br.p + br.pos + len);
            br.pos += len;
            continue;
        }
        if(type == 3)throw std::runtime_error("inflate: bad block type");
        Huffman lit, dist;
        uint8_t lens[320];
        if(type == 1){
            for(int i = 0; i < 288; i++)lens[i] = i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8;
            lit.build(lens, 288);
            for(int i = 0; i < 30; i++)lens[i] = 5;
            dist.build(lens, 30);
        } else{
            const int nl = (int)br.bits(5) + 257, nd = (int)br.bits(5) + 1, nc = (int)br.bits(4) + 4;
            static const uint8_t order[19] = { 16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15 };
            uint8_t cl[19] = {};
            for(int i = 0; i < nc; i++)cl[order[i]] = (uint8_t)br.bits(3);
            Huffman ch; ch.build(cl, 19);
            for(int i = 0; i < nl + nd; ){
                const int s = ch.decode(br);
                if(s < 16){ lens[i++] = (uint8_t)s; continue; }
                int rep; uint8_t v = 0;
                if(s == 16){ if(i == 0)throw std::runtime_error("inflate: repeat without a previous length"); v = lens[i - 1]; rep = 3 + (int)br.bits(2); }
                else if(s == 17)rep = 3 + (int)br.bits(3);
                else rep = 11 + (int)br.bits(7);
                if(i + rep > nl + nd)throw std::runtime_error("inflate: too many code lengths");
                while(rep--)lens[i++] = v;
            }
            lit.build(lens, nl);
            dist.build(lens + nl, nd);
        }
        for(;;){
            const int s = lit.decode(br);
            if(s < 256){ out.push_back((uint8_t)s); continue; }
            if(s == 256)break;
            if(s > 285)throw std::runtime_error("inflate: bad length symbol");
            const int len = lbase[s - 257] + (int)br.bits(lext[s - 257]);
            const int ds = dist.decode(br);
            if(ds > 29)throw std::runtime_error("inflate: bad distance symbol");
            const size_t d = dbase[ds] + br.bits(dext[ds]);
            if(d > out.size())throw std::runtime_error("inflate: distance beyond the output");
            const size_t from = out.size() - d;
            for(int k = 0; k < len; k++)out.push_back(out[from + k]);
        }
    }
    return out;
}

// ---- PNG ---------------------------------------------------------------------------------------------------------
inline ImageU8 decode_png(const std::vector<uint8_t>& f){
    auto be32 = [&](size_t o){ return (uint32_t)f[o] << 24 | (uint32_t)f[o + 1] << 16 | (uint32_t)f[o + 2] << 8 | f[o + 3]; };
    if(f.size() < 33)throw std::runtime_error("png: truncated");
    uint32_t w = 0, h = 0; int depth = 0, ctype = 0, interlace = 0;
    std::vector<uint8_t> idat, plte, trns;
    for(size_t o = 8; o + 12 <= f.size(); ){
        const uint32_t len = be32(o);
        const std::string type((const char*)&f[o + 4], 4);
        if(o + 12 + len > f.size())throw std::runtime_error("png: truncated chunk");
        const uint8_t* d = &f[o + 8];
        if(type == "IHDR"){ w = be32(o + 8); h = be32(o + 12); depth = d[8]; ctype = d[9]; interlace = d[12]; }
        else if(type == "IDAT")idat.insert(idat.end(), d, d + len);
        else if(type == "PLTE")plte.assign(d, d + len);
        else if(type == "tRNS")trns.assign(d, d + len);
        else if(type == "IEND")break;
        o += 12 + len;
    }
    if(w == 0 || h == 0 || interlace)throw std::runtime_error("png: interlaced or empty images are not supported");
    if(!(depth == 8 || depth == 16 || (ctype == 3 && (depth == 1 || depth == 2 || depth == 4)) || (ctype == 0 && depth < 8)))throw std::runtime_error("png: unsupported bit depth");
    const int samples = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : 4;
    const size_t bpp = std::max<size_t>(1, (size_t)samples * depth / 8), stride = ((size_t)w * samples * depth + 7) / 8;
    std::vector<uint8_t> raw = inflate_zlib(idat.data(), idat.size());
    if(raw.size() < (stride + 1) * h)throw std::runtime_error("png: image data too short");
    std::vector<uint8_t> img(stride * h);
    for(uint32_t y = 0; y < h; y++){
        const uint8_t ft = raw[y * (stride + 1)];
        const uint8_t* in = &raw[y * (stride + 1) + 1];
        uint8_t* out = &img[y * stride];
        const uint8_t* up = y ? out - stride : nullptr;
        for(size_t x = 0; x < stride; x++){
            const int a = x >= bpp ? out[x - bpp] : 0, b = up ? up[x] : 0, c = (up && x >= bpp) ? up[x - bpp] : 0;
            int pred = 0;
            switch(ft){
            case 0: break;
            case 1: pred = a; break;
            case 2: pred = b; break;
            case 3: pred = (a + b) >> 1; break;
            case 4: { const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c); pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); break; }
            default: throw std::runtime_error("png: bad filter type");
            }
            out[x] = (uint8_t)(in[x] + pred);
        }
    }
    ImageU8 r; r.width = (int)w; r.height = (int)h;
    auto sample = [&](uint32_t y, size_t idx) -> int{            // idx-th sample of row y, scaled to 8 bits
        const uint8_t* row = &img[y * stride];
        if(depth == 8)return row[idx];
        if(depth == 16)return row[2 * idx];
        const int per = 8 / depth, v = (row[idx / per] >> ((per - 1 - idx % per) * depth)) & ((1 << depth) - 1);
        return ctype == 3 ? v : v * 255 / ((1 << depth) - 1);
    };
    if(ctype == 3){
        r.channels = trns.empty() ? 3 : 4;
        r.data.resize((size_t)w * h * r.channels);
        for(uint32_t y = 0; y < h; y++)for(uint32_t x = 0; x < w; x++){
            const int i = sample(y, x);
            if((size_t)3 * i + 2 >= plte.size())throw std::runtime_error("png: palette index out of range");
            uint8_t* o = &r.data[((size_t)y * w + x) * r.channels];
            o[0] = plte[3 * i]; o[1] = plte[3 * i + 1]; o[2] = plte[3 * i + 2];
            if(r.channels == 4)o[3] = (size_t)i < trns.size() ? trns[i] : 255;
        }
    } else{
        r.channels = samples;
        r.data.resize((size_t)w * h * samples);
        for(uint32_t y = 0; y < h; y++)for(size_t i = 0; i < (size_t)w * samples; i++)r.data[(size_t)y * w * samples + i] = (uint8_t)sample(y, i);
    }
    return r;
}

// ---- PPM / PGM (binary), PFM ----------------------------------------------------------------------------------------
namespace detail {
inline bool pnm_token(const std::vector<uint8_t>& f, size_t& o, std::string& tok){
    for(;;){
        while(o < f.size() && std::isspace(f[o]))o++;
        if(o < f.size() && f[o] == '#'){ while(o < f.size() && f[o] != '\n')o++; continue; }
        break;
    }
    tok.clear();
    while(o < f.size() && !std::isspace(f[o]))tok.push_back((char)f[o++]);
    return !tok.empty();
}
}
inline ImageU8 decode_pnm(const std::vector<uint8_t>& f){
    size_t o = 0; std::string magic, a, b, c;
    if(!detail::pnm_token(f, o, magic) || !detail::pnm_token(f, o, a) || !detail::pnm_token(f, o, b) || !detail::pnm_token(f, o, c))throw std::runtime_error("pnm: bad header");
    ImageU8 r; r.width = std::stoi(a); r.height = std::stoi(b); r.channels = magic == "P6" ? 3 : 1;
    const int maxv = std::stoi(c);
    if((magic != "P6" && magic != "P5") || maxv <= 0 || maxv > 255)throw std::runtime_error("pnm: only binary P5 / P6 with maxval <= 255");
    o++;        // the single whitespace after maxval
    const size_t n = (size_t)r.width * r.height * r.channels;
    if(o + n > f.size())throw std::runtime_error("pnm: truncated");
    r.data.assign(f.begin() + (long)o, f.begin() + (long)(o + n));
    if(maxv != 255)for(uint8_t& v : r.data)v = (uint8_t)(v * 255 / maxv);
    return r;
}
inline ImageF32 decode_pfm(const std::vector<uint8_t>& f){
    size_t o = 0; std::string magic, a, b, c;
    if(!detail::pnm_token(f, o, magic) || !detail::pnm_token(f, o, a) || !detail::pnm_token(f, o, b) || !detail::pnm_token(f, o, c))throw std::runtime_error("pfm: bad header");
    ImageF32 r; r.width = std::stoi(a); r.height = std::stoi(b); r.channels = magic == "PF" ? 3 : 1;
    if(magic != "PF" && magic != "Pf")throw std::runtime_error("pfm: bad magic");
    const bool little = std::stof(c) < 0;
    o++;
    const size_t n = (size_t)r.width * r.height * r.channels;
    if(o + 4 * n > f.size())throw std::runtime_error("pfm: truncated");
    r.data.resize(n);
    for(int y = 0; y < r.height; y++)for(size_t i = 0; i < (size_t)r.width * r.channels; i++){       // PFM stores the bottom row first
        const uint8_t* p = &f[o + 4 * ((size_t)(r.height - 1 - y) * r.width * r.channels + i)];
        uint32_t u = little ? (uint32_t)p[0] | p[1] << 8 | p[2] << 16 | (uint32_t)p[3] << 24 : (uint32_t)p[3] | p[2] << 8 | p[1] << 16 | (uint32_t)p[0] << 24;
        std::memcpy(&r.data[(size_t)y * r.width * r.channels + i], &u, 4);
    }
    return r;
}

// ---- BMP (uncompressed), TGA ------------------------------------------------------------------------------------------
inline ImageU8 decode_bmp(const std::vector<uint8_t>& f){
    auto le32 = [&](size_t o){ return (uint32_t)f[o] | f[o + 1] << 8 | f[o + 2] << 16 | (uint32_t)f[o + 3] << 24; };
    if(f.size() < 54)throw std::runtime_error("bmp: truncated");
    const uint32_t off = le32(10);
    const int32_t w = (int32_t)le32(18), hs = (int32_t)le32(22);
    const int bpp = f[28] | f[29] << 8, comp = (int)le32(30);
    if((bpp != 24 && bpp != 32) || (comp != 0 && comp != 3) || w <= 0 || hs == 0)throw std::runtime_error("bmp: only uncompressed 24 / 32 bit");
    const int h = std::abs(hs);
    ImageU8 r; r.width = w; r.height = h; r.channels = bpp / 8;
    const size_t stride = ((size_t)w * r.channels + 3) & ~(size_t)3;
    if(off + stride * h > f.size())throw std::runtime_error("bmp: truncated");
    r.data.resize((size_t)w * h * r.channels);
    for(int y = 0; y < h; y++){
        const uint8_t* row = &f[off + stride * (size_t)(hs > 0 ? h - 1 - y : y)];
        for(int x = 0; x < w; x++){
            uint8_t* o = &r.data[((size_t)y * w + x) * r.channels];
            o[0] = row[x * r.channels + 2]; o[1] = row[x * r.channels + 1]; o[2] = row[x * r.channels];
            if(r.channels == 4)o[3] = row[x * 4 + 3];
        }
    }
    return r;
}
inline ImageU8 decode_tga(const std::vector<uint8_t>& f){
    if(f.size() < 18)throw std::runtime_error("tga: truncated");
    const int idlen = f[0], cmap = f[1], type = f[2], w = f[12] | f[13] << 8, h = f[14] | f[15] << 8, bpp = f[16], desc = f[17];
    if(cmap != 0 || !(type == 2 || type == 3 || type == 10 || type == 11) || !(bpp == 8 || bpp == 24 || bpp == 32) || w <= 0 || h <= 0)throw std::runtime_error("tga: unsupported variant");
    const int ch = bpp / 8;
    ImageU8 r; r.width = w; r.height = h; r.channels = ch;
    r.data.resize((size_t)w * h * ch);
    size_t o = 18 + (size_t)idlen, px = 0;
    const size_t total = (size_t)w * h;
    std::vector<uint8_t> lin(total * ch);
    auto need = [&](size_t k){ if(o + k > f.size())throw std::runtime_error("tga: truncated"); };
    if(type < 8){ need(total * ch); std::memcpy(lin.data(), &f[o], total * ch); }
    else while(px < total){
        need(1);
        const int hd = f[o++], cnt = (hd & 127) + 1;
        if(px + cnt > total)throw std::runtime_error("tga: run beyond the image");
        if(hd & 128){ need(ch); for(int k = 0; k < cnt; k++)std::memcpy(&lin[(px + k) * ch], &f[o], ch); o += ch; }
        else{ need((size_t)cnt * ch); std::memcpy(&lin[px * ch], &f[o], (size_t)cnt * ch); o += (size_t)cnt * ch; }
        px += cnt;
    }
    const bool top = (desc & 0x20) != 0;
    for(int y = 0; y < h; y++)for(int x = 0; x < w; x++){
        const uint8_t* s = &lin[((size_t)(top ? y : h - 1 - y) * w + x) * ch];
        uint8_t* d = &r.data[((size_t)y * w + x) * ch];
        if(ch == 1)d[0] = s[0];
        else{ d[0] = s[2]; d[1] = s[1]; d[2] = s[0]; if(ch == 4)d[3] = s[3]; }
    }
    return r;
}

// ---- Radiance .hdr (RGBE) ---------------------------------------------------------------------------------------------
inline ImageF32 decode_hdr(const std::vector<uint8_t>& f){
    size_t o = 0;
    auto line = [&](){ std::string s; while(o < f.size() && f[o] != '\n')s.push_back((char)f[o++]); o++; return s; };
    const std::string magic = line();
    if(magic != "#?RADIANCE" && magic != "#?RGBE")throw std::runtime_error("hdr: bad signature");
    for(;;){ if(o >= f.size())throw std::runtime_error("hdr: truncated header"); const std::string l = line(); if(l.empty())break; }
    const std::string res = line();
    int w = 0, h = 0;
    if(std::sscanf(res.c_str(), "-Y %d +X %d", &h, &w) != 2 || w <= 0 || h <= 0)throw std::runtime_error("hdr: only -Y +X orientation");
    ImageF32 r; r.width = w; r.height = h; r.channels = 3;
    r.data.resize((size_t)w * h * 3);
    std::vector<uint8_t> scan((size_t)w * 4);
    auto need = [&](size_t k){ if(o + k > f.size())throw std::runtime_error("hdr: truncated"); };
    for(int y = 0; y < h; y++){
        need(4);
        if(w >= 8 && w < 32768 && f[o] == 2 && f[o + 1] == 2 && !(f[o + 2] & 0x80)){
            if(((int)f[o + 2] << 8 | f[o + 3]) != w)throw std::runtime_error("hdr: scanline width mismatch");
            o += 4;
            for(int c = 0; c < 4; c++)for(int x = 0; x < w; ){
                need(1);
                int cnt = f[o++];
                if(cnt > 128){ cnt -= 128; need(1); if(x + cnt > w)throw std::runtime_error("hdr: bad run"); const uint8_t v = f[o++]; while(cnt--)scan[(size_t)(x++) * 4 + c] = v; }
                else{ if(cnt == 0 || x + cnt > w)throw std::runtime_error("hdr: bad run"); need((size_t)cnt); while(cnt--)scan[(size_t)(x++) * 4 + c] = f[o++]; }
            }
        } else{ need((size_t)w * 4); std::memcpy(scan.data(), &f[o], (size_t)w * 4); o += (size_t)w * 4; }
        for(int x = 0; x < w; x++){
            const uint8_t* p = &scan[(size_t)x * 4];
            float* d = &r.data[((size_t)y * w + x) * 3];
            if(p[3] == 0){ d[0] = d[1] = d[2] = 0.0f; continue; }
            const float s = std::ldexp(1.0f, (int)p[3] - (128 + 8));        // stb_image's convention: mantissa / 256 * 2^(e - 128)
            d[0] = p[0] * s; d[1] = p[1] * s; d[2] = p[2] * s;
        }
    }
    return r;
}

// ---- front ends: what stbi_load / stbi_loadf return for the reference's texture constructors ---------------------------
inline ImageU8 load_u8(const std::string& path){
    const std::vector<uint8_t> f = read_file(path);
    if(f.size() >= 8 && !std::memcmp(f.data(), "\x89PNG\r\n\x1a\n", 8))return decode_png(f);
    if(f.size() >= 2 && f[0] == 'P' && (f[1] == '5' || f[1] == '6'))return decode_pnm(f);
    if(f.size() >= 2 && f[0] == 'B' && f[1] == 'M')return decode_bmp(f);
    if(f.size() >= 3 && f[0] == 0xFF && f[1] == 0xD8)throw std::runtime_error("JPEG decoding is not implemented (" + path + "): convert the texture to PNG");
    const size_t dot = path.rfind('.');
    std::string ext = dot == std::string::npos ? "" : path.substr(dot + 1);
    for(char& c : ext)c = (char)std::tolower(c);
    if(ext == "tga")return decode_tga(f);
    throw std::runtime_error("unknown image format: " + path);
}
inline ImageF32 load_f32(const std::string& path){
    const std::vector<uint8_t> f = read_file(path);
    if(f.size() >= 2 && f[0] == '#' && f[1] == '?')return decode_hdr(f);
    if(f.size() >= 2 && f[0] == 'P' && (f[1] == 'F' || f[1] == 'f'))return decode_pfm(f);
    // an LDR file through stbi_loadf: (v / 255) ^ 2.2 on the colour channels, alpha linear
    const ImageU8 u = load_u8(path);
    ImageF32 r; r.width = u.width; r.height = u.height; r.channels = u.channels;
    r.data.resize(u.data.size());
    const int color = (u.channels & 1) ? u.channels : u.channels - 1;
    for(size_t i = 0; i < u.data.size(); i++)
        r.data[i] = (int)(i % (size_t)u.channels) < color ? std::pow(u.data[i] / 255.0f, 2.2f) : u.data[i] / 255.0f;
    return r;
}

// ---- baseline JPEG encoder (ITU T.81 Annex K tables, JFIF 1.01, 4:4:4) -------------------------------------------------
namespace detail {
struct JpegBits {
    std::vector<uint8_t>& out; uint32_t acc = 0; int cnt = 0;
    void put(uint32_t code, int len){
        acc = (acc << len) | (code & ((1u << len) - 1u)); cnt += len;
        while(cnt >= 8){ const uint8_t b = (uint8_t)(acc >> (cnt - 8)); out.push_back(b); if(b == 0xFF)out.push_back(0); cnt -= 8; }
    }
    void flush(){ if(cnt)put(0x7F, 8 - cnt); }
};
inline void jpeg_codes(const uint8_t* bits, const uint8_t* vals, uint16_t* code, uint8_t* size){
    uint16_t c = 0; int k = 0;
    for(int l = 1; l <= 16; l++){ for(int i = 0; i < bits[l - 1]; i++){ code[vals[k]] = c++; size[vals[k]] = (uint8_t)l; k++; } c <<= 1; }
}
}
inline void write_jpg(const std::string& path, int w, int h, const uint8_t* rgb /* row 0 = top */, int quality){
    using namespace detail;
    static const uint8_t zig[64] = { 0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22,
                                     15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63 };
    static const uint8_t qy[64] = { 16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55, 14, 13, 16, 24, 40, 57, 69, 56, 14, 17, 22, 29, 51, 87, 80, 62, 18, 22, 37, 56, 68, 109, 103, 77,
                                    24, 35, 55, 64, 81, 104, 113, 92, 49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99 };
    static const uint8_t qc[64] = { 17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99, 99, 99, 47, 66, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
                                    99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99 };
    static const uint8_t dc_l_bits[16] = { 0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0 }, dc_c_bits[16] = { 0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0 };
    static const uint8_t dc_vals[12] = { 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11 };
    static const uint8_t ac_l_bits[16] = { 0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 0x7d }, ac_c_bits[16] = { 0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 0x77 };
    static const uint8_t ac_l_vals[162] = { 0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07, 0x22, 0x71, 0x14, 0x32, 0x81, 0x91, 0xa1, 0x08, 0x23, 0x42, 0xb1, 0xc1,
        0x15, 0x52, 0xd1, 0xf0, 0x24, 0x33, 0x62, 0x72, 0x82, 0x09, 0x0a, 0x16, 0x17, 0x18, 0x19, 0x1a, 0x25, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x34, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46,
        0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x83, 0x84, 0x85, 0x86, 0x87,
        0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3, 0xc4,
        0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda, 0xe1, 0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf1, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8,
        0xf9, 0xfa };
    static const uint8_t ac_c_vals[162] = { 0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71, 0x13, 0x22, 0x32, 0x81, 0x08, 0x14, 0x42, 0x91, 0xa1, 0xb1, 0xc1, 0x09,
        0x23, 0x33, 0x52, 0xf0, 0x15, 0x62, 0x72, 0xd1, 0x0a, 0x16, 0x24, 0x34, 0xe1, 0x25, 0xf1, 0x17, 0x18, 0x19, 0x1a, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45,
        0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x82, 0x83, 0x84, 0x85,
        0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2,
        0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda, 0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8,
        0xf9, 0xfa };
    quality = quality < 1 ? 1 : quality > 100 ? 100 : quality;
    const int scale = quality < 50 ? 5000 / quality : 200 - quality * 2;
    uint8_t qt[2][64];
    for(int i = 0; i < 64; i++){
        qt[0][i] = (uint8_t)std::min(255, std::max(1, (qy[i] * scale + 50) / 100));
        qt[1][i] = (uint8_t)std::min(255, std::max(1, (qc[i] * scale + 50) / 100));
    }
    uint16_t dcc[2][12], acc_[2][256]; uint8_t dcs[2][12], acs[2][256];
    std::memset(acs, 0, sizeof acs);
    jpeg_codes(dc_l_bits, dc_vals, dcc[0], dcs[0]); jpeg_codes(dc_c_bits, dc_vals, dcc[1], dcs[1]);
    jpeg_codes(ac_l_bits, ac_l_vals, acc_[0], acs[0]); jpeg_codes(ac_c_bits, ac_c_vals, acc_[1], acs[1]);
    std::vector<uint8_t> o;
    auto u16 = [&](int v){ o.push_back((uint8_t)(v >> 8)); o.push_back((uint8_t)v); };
    o.push_back(0xFF); o.push_back(0xD8);
    o.push_back(0xFF); o.push_back(0xE0); u16(16); for(char c : { 'J', 'F', 'I', 'F', '\0' })o.push_back((uint8_t)c); o.push_back(1); o.push_back(1); o.push_back(0); u16(1); u16(1); o.push_back(0); o.push_back(0);
    for(int t = 0; t < 2; t++){ o.push_back(0xFF); o.push_back(0xDB); u16(67); o.push_back((uint8_t)t); for(int i = 0; i < 64; i++)o.push_back(qt[t][zig[i]]); }
    o.push_back(0xFF); o.push_back(0xC0); u16(17); o.push_back(8); u16(h); u16(w); o.push_back(3);
    for(int c = 0; c < 3; c++){ o.push_back((uint8_t)(c + 1)); o.push_back(0x11); o.push_back(c ? 1 : 0); }
    auto dht = [&](int cls_id, const uint8_t* bits, const uint8_t* vals, int n){ o.push_back(0xFF); o.push_back(0xC4); u16(19 + n); o.push_back((uint8_t)cls_id); o.insert(o.end(), bits, bits + 16); o.insert(o.end(), vals, vals + n); };
    dht(0x00, dc_l_bits, dc_vals, 12); dht(0x10, ac_l_bits, ac_l_vals, 162); dht(0x01, dc_c_bits, dc_vals, 12); dht(0x11, ac_c_bits, ac_c_vals, 162);
    o.push_back(0xFF); o.push_back(0xDA); u16(12); o.push_back(3);
    for(int c = 0; c < 3; c++){ o.push_back((uint8_t)(c + 1)); o.push_back(c ? 0x11 : 0x00); }
    o.push_back(0); o.push_back(63); o.push_back(0);
    JpegBits bw { o };
    double cs[8][8];
    for(int u = 0; u < 8; u++)for(int x = 0; x < 8; x++)cs[u][x] = std::cos((2 * x + 1) * u * 3.14159265358979323846 / 16.0) * (u == 0 ? std::sqrt(0.125) : 0.5);
    int pred[3] = { 0, 0, 0 };
    for(int by = 0; by < h; by += 8)for(int bx = 0; bx < w; bx += 8){
        double blk[3][64];
        for(int y = 0; y < 8; y++)for(int x = 0; x < 8; x++){
            const uint8_t* p = rgb + 3 * ((size_t)std::min(h - 1, by + y) * w + std::min(w - 1, bx + x));
            const double r = p[0], g = p[1], b = p[2];
            blk[0][y * 8 + x] = 0.299 * r + 0.587 * g + 0.114 * b - 128.0;
            blk[1][y * 8 + x] = -0.168736 * r - 0.331264 * g + 0.5 * b;
            blk[2][y * 8 + x] = 0.5 * r - 0.418688 * g - 0.081312 * b;
        }
        for(int c = 0; c < 3; c++){
            const int t = c ? 1 : 0;
            double tmp[64], dct[64];
            for(int y = 0; y < 8; y++)for(int u = 0; u < 8; u++){ double s = 0; for(int x = 0; x < 8; x++)s += blk[c][y * 8 + x] * cs[u][x]; tmp[y * 8 + u] = s; }
            for(int v = 0; v < 8; v++)for(int u = 0; u < 8; u++){ double s = 0; for(int y = 0; y < 8; y++)s += tmp[y * 8 + u] * cs[v][y]; dct[v * 8 + u] = s; }
            int q[64];
            for(int i = 0; i < 64; i++)q[i] = (int)std::lround(dct[zig[i]] / qt[t][zig[i]]);
            auto cat = [](int v){ int a = std::abs(v), n = 0; while(a){ n++; a >>= 1; } return n; };
            auto amp = [](int v, int n){ return (uint32_t)(v < 0 ? v + (1 << n) - 1 : v); };
            const int diff = q[0] - pred[c]; pred[c] = q[0];
            int n = cat(diff);
            bw.put(dcc[t][n], dcs[t][n]);
            if(n)bw.put(amp(diff, n), n);
            int run = 0, lastnz = 63;
            while(lastnz > 0 && q[lastnz] == 0)lastnz--;
            for(int i = 1; i <= lastnz; i++){
                if(q[i] == 0){ run++; continue; }
                while(run > 15){ bw.put(acc_[t][0xF0], acs[t][0xF0]); run -= 16; }
                n = cat(q[i]);
                bw.put(acc_[t][run << 4 | n], acs[t][run << 4 | n]);
                bw.put(amp(q[i], n), n);
                run = 0;
            }
            if(lastnz < 63)bw.put(acc_[t][0x00], acs[t][0x00]);
        }
    }
    bw.flush();
    o.push_back(0xFF); o.push_back(0xD9);
    std::ofstream f(path, std::ios::binary);
    if(!f)throw std::runtime_error("write_jpg: cannot open " + path);
    f.write((const char*)o.data(), (std::streamsize)o.size());
}

} // namespace ptb_io
