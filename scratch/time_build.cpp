#include "../pathtracing_b200/csrc/bvh_build.hpp"
#include <chrono>
#include <cstdio>
#include <random>
using namespace ptb;
int main(int argc,char**argv){
    uint32_t n = argc>1? atoi(argv[1]):1000000; unsigned T = argc>2? atoi(argv[2]):0;
    std::mt19937 g(42); std::uniform_real_distribution<float> U(0,1);
    std::vector<Box> boxes(n);
    for(auto&b:boxes){ float c[3]={U(g),U(g),U(g)}; b=box_empty(); for(int v=0;v<3;v++){float p[3]; for(int k=0;k<3;k++)p[k]=c[k]+0.01f*(U(g)-0.5f); box_grow(b,p);} }
    auto t0=std::chrono::steady_clock::now();
    BinaryBvh bb = build_binary_bvh(boxes.data(), n);
    auto t1=std::chrono::steady_clock::now();
    FlatBvh f4 = flatten_bvh(bb,4,T);
    auto t2=std::chrono::steady_clock::now();
    FlatBvh f2 = flatten_bvh(bb,2,T);
    auto t3=std::chrono::steady_clock::now();
    FlatBvh f8 = flatten_bvh(bb,8,T);
    auto t4=std::chrono::steady_clock::now();
    auto ms=[](auto a,auto b){return std::chrono::duration<double,std::milli>(b-a).count();};
    printf("n=%u nodes=%zu build %.1f ms flatten4 %.1f (n4=%zu) flatten2 %.1f flatten8 %.1f\n",n,bb.nodes.size(),ms(t0,t1),ms(t1,t2),f4.n4.size(),ms(t2,t3),ms(t3,t4));
}
