// tests/host/imageio_tool.cpp — TEST TOOL: drives pathtracing_b200/host/ImageIO.hpp and the Film writers of PathTracing.hpp
// from the command line so that tests/test_imageio.py can compare them with Pillow / OpenCV and with the reference's writers.
//   decode8  <file> <out.raw>        : load_u8  -> "W H C\n" + bytes
//   decodef  <file> <out.raw>        : load_f32 -> "W H C\n" + float32
//   film     <in.f32> W H <out_noext> <ppm|png|jpg> [quality] : Film with these pixel values (row 0 = bottom) -> writer
//   texture8 <file> gamma            : ImageTexture(file, gamma) constructs (prints W H C of what would be uploaded)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include "PathTracing.hpp"

int main(int argc, char** argv){
    try{
        const std::string cmd = argc > 1 ? argv[1] : "";
        if(cmd == "decode8" && argc == 4){
            const ptb_io::ImageU8 im = ptb_io::load_u8(argv[2]);
            std::ofstream o(argv[3], std::ios::binary);
            o << im.width << " " << im.height << " " << im.channels << "\n";
            o.write((const char*)im.data.data(), (std::streamsize)im.data.size());
            return 0;
        }
        if(cmd == "decodef" && argc == 4){
            const ptb_io::ImageF32 im = ptb_io::load_f32(argv[2]);
            std::ofstream o(argv[3], std::ios::binary);
            o << im.width << " " << im.height << " " << im.channels << "\n";
            o.write((const char*)im.data.data(), (std::streamsize)(im.data.size() * 4));
            return 0;
        }
        if(cmd == "film" && argc >= 7){
            const int W = std::atoi(argv[3]), H = std::atoi(argv[4]);
            std::vector<float> rgb((size_t)W * H * 3);
            std::ifstream in(argv[2], std::ios::binary);
            in.read((char*)rgb.data(), (std::streamsize)(rgb.size() * 4));
            auto film = std::make_shared<Film>(glm::ivec2(W, H), std::make_shared<BoxFilter>());
            std::vector<double> rgbw((size_t)W * H * 4);
            for(size_t i = 0; i < (size_t)W * H; i++){ rgbw[4 * i] = rgb[3 * i]; rgbw[4 * i + 1] = rgb[3 * i + 1]; rgbw[4 * i + 2] = rgb[3 * i + 2]; rgbw[4 * i + 3] = 1.0; }
            film->Merge(rgbw);
            film->setOutputPath(argv[5]);
            const std::string fmt = argv[6];
            if(fmt == "ppm")film->WritePPM("unused");
            else if(fmt == "png")film->WritePNG("unused");
            else film->WriteJPG("unused", argc > 7 ? std::atoi(argv[7]) : 90);
            return 0;
        }
        if(cmd == "texture8" && argc == 4){
            ImageTexture t(argv[2], std::atoi(argv[3]) != 0);
            std::printf("channels %d\n", t.Channels());
            return 0;
        }
        if(cmd == "sky"){
            // an arbitrary sky function goes through the bake: report its error
            FunctionInfiniteLight sun([](const Ray& r){ const float s = std::max(0.0f, r.dir.y); return glm::vec3(0.2f + 3.0f * s * s, 0.3f + 0.5f * s, 0.6f + 0.2f * r.dir.x); });
            const auto e = sun.BakeError();
            std::printf("baked %d max_rel %.6f mean_rel %.6f\n", sun.Baked() ? 1 : 0, e.max_rel, e.mean_rel);
            FunctionInfiniteLight grad(GradientSky{});
            std::printf("gradient baked %d\n", grad.Baked() ? 1 : 0);
            return 0;
        }
        std::fprintf(stderr, "usage: see the head of imageio_tool.cpp\n");
        return 2;
    } catch(const std::exception& e){
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
}
