// Mutation fuzz of the vsp_codec_* blob readers (truncation, bit flips, forged counts, trailing bytes, swaps) over exact-size
// heap copies, for an AddressSanitizer / UBSan build of csrc/codec.cpp: tests/test_codec.py builds and runs it.
// usage: codec_fuzz <iterations> <directory with pub.bin svk.bin vk.bin ct.bin pk.bin pi.bin>
#include <unistd.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>
#include "vsp_b200.h"
static std::vector<uint8_t> load(const char* f){FILE*p=fopen(f,"rb");std::vector<uint8_t> v;int c;while((c=fgetc(p))!=EOF)v.push_back((uint8_t)c);fclose(p);return v;}
static std::mt19937_64 rng(12345);
static std::vector<uint8_t> mutate(const std::vector<uint8_t>& in){
    std::vector<uint8_t> v=in; int kind=rng()%5;
    if(kind==0&&!v.empty()) v.resize(rng()%v.size());
    else if(kind==1){ for(int i=0;i<1+(int)(rng()%4);++i) if(!v.empty()) v[rng()%v.size()]^=(uint8_t)(1u<<(rng()%8)); }
    else if(kind==2&&v.size()>8){ size_t pos=rng()%(v.size()-8); uint64_t x=rng(); if(rng()%2) x%=1000; for(int i=0;i<8;++i) v[pos+i]=(uint8_t)(x>>(8*(7-i))); }
    else if(kind==3){ for(int i=0;i<(int)(rng()%64);++i) v.push_back((uint8_t)rng()); }
    else if(kind==4&&v.size()>16){ size_t a=rng()%v.size(), b=rng()%v.size(); std::swap(v[a],v[b]); }
    // exact-size heap copy so ASAN sees any over-read
    return std::vector<uint8_t>(v.begin(), v.end());
}
int main(int argc,char**argv){
    const int iters = argc > 1 ? atoi(argv[1]) : 300;
    if (argc > 2 && chdir(argv[2]) != 0) return 2;
    auto pub=load("pub.bin"),svk=load("svk.bin"),vk=load("vk.bin"),ct=load("ct.bin"),pk=load("pk.bin"),pi=load("pi.bin");
    long ok=0,bad=0;
    for(int it=0;it<iters;++it){
        { auto b=mutate(pub); uint8_t* d=(uint8_t*)malloc(b.size()?b.size():1); if(!b.empty()) memcpy(d,b.data(),b.size()); uint64_t n=0;
          if(vsp_codec_saver_pubkey_read(d,b.size(),&n,nullptr)==VSP_OK && n<100000){
            std::vector<uint64_t> a(12),b1(12),c(12),ds(12*n+12),t1(12*n+12),t2(24*(n+1));
            vsp_saver_pubkey_t p; p.msg_size=n; p.delta_g1=a.data(); p.delta_sum_s_g1=b1.data(); p.gamma_inverse_sum_s_g1=c.data(); p.delta_s_g1=ds.data(); p.t_g1=t1.data(); p.t_g2=t2.data();
            if(vsp_codec_saver_pubkey_read(d,b.size(),nullptr,&p)==VSP_OK) ++ok; else ++bad; } else ++bad; free(d);}
        { auto b=mutate(svk); uint8_t* d=(uint8_t*)malloc(b.size()?b.size():1); if(!b.empty()) memcpy(d,b.data(),b.size()); uint64_t n=0;
          if(vsp_codec_saver_vk_read(d,b.size(),&n,nullptr,nullptr,nullptr,0)==VSP_OK && n<100000){
            std::vector<uint64_t> a(24),s1(24*n+24),s2(24*n+24);
            if(vsp_codec_saver_vk_read(d,b.size(),nullptr,a.data(),s1.data(),s2.data(),n)==VSP_OK) ++ok; else ++bad; } else ++bad; free(d);}
        { auto b=mutate(vk); uint8_t* d=(uint8_t*)malloc(b.size()?b.size():1); if(!b.empty()) memcpy(d,b.data(),b.size()); uint64_t n=0;
          if(vsp_codec_groth16_vk_read(d,b.size(),&n,nullptr,nullptr,nullptr,nullptr,nullptr,nullptr,0)==VSP_OK && n<1000000){
            std::vector<uint64_t> gt(72),g2a(24),g2b(24),g1a(12),g1b(12),abc(12*n);
            if(vsp_codec_groth16_vk_read(d,b.size(),nullptr,gt.data(),g2a.data(),g2b.data(),g1a.data(),g1b.data(),abc.data(),n)==VSP_OK) ++ok; else ++bad; } else ++bad; free(d);}
        { auto b=mutate(ct); uint8_t* d=(uint8_t*)malloc(b.size()?b.size():1); if(!b.empty()) memcpy(d,b.data(),b.size()); uint64_t n=0;
          if(vsp_codec_ciphertext_read(d,b.size(),&n,nullptr,0)==VSP_OK){ std::vector<uint64_t> o(12*n+12);
            if(vsp_codec_ciphertext_read(d,b.size(),&n,o.data(),n)==VSP_OK) ++ok; else ++bad; } else ++bad; free(d);}
        { auto b=mutate(pi); uint8_t* d=(uint8_t*)malloc(b.size()?b.size():1); if(!b.empty()) memcpy(d,b.data(),b.size()); uint64_t n=0;
          if(b.size()>=8 && vsp_codec_primary_input_read(d,b.size(),&n,nullptr,0)==VSP_OK){ std::vector<uint64_t> o(4*n+4);
            if(vsp_codec_primary_input_read(d,b.size(),&n,o.data(),n)==VSP_OK) ++ok; else ++bad; } else ++bad; free(d);}
        if(it%4==0){ auto b=mutate(pk); uint8_t* d=(uint8_t*)malloc(b.size()?b.size():1); if(!b.empty()) memcpy(d,b.data(),b.size()); vsp_fast_pk_sizes_t sz;
          if(vsp_codec_fast_pk_sizes(d,b.size(),&sz)==VSP_OK){
            uint64_t nc=sz.num_constraints;
            std::vector<uint64_t> ar(nc+1),br(nc+1),cr(nc+1),av(4*sz.nnz_a+4),bv(4*sz.nnz_b+4),cv(4*sz.nnz_c+4);
            std::vector<uint32_t> ac(sz.nnz_a+1),bc(sz.nnz_b+1),cc(sz.nnz_c+1),bi(sz.B_count+1);
            std::vector<uint64_t> p1(12),p2(12),p3(24),p4(12),p5(24),A(12*sz.A_count+12),B2(24*sz.B_count+24),B1(12*sz.B_count+12),H(12*sz.H_count+12),L(12*sz.L_count+12);
            vsp_proving_key_t k; memset(&k,0,sizeof(k));
            k.cs.a_row=ar.data();k.cs.b_row=br.data();k.cs.c_row=cr.data();k.cs.a_col=ac.data();k.cs.b_col=bc.data();k.cs.c_col=cc.data();k.cs.a_val=av.data();k.cs.b_val=bv.data();k.cs.c_val=cv.data();
            k.alpha_g1=p1.data();k.beta_g1=p2.data();k.beta_g2=p3.data();k.delta_g1=p4.data();k.delta_g2=p5.data();k.A_query=A.data();k.B_idx=bi.data();k.B_g2=B2.data();k.B_g1=B1.data();k.H_query=H.data();k.L_query=L.data();
            if(vsp_codec_fast_pk_read(d,b.size(),&k)==VSP_OK) ++ok; else ++bad; } else ++bad; free(d);}
    }
    // bit arrays / tree
    for(int it=0;it<iters;++it){ size_t len=rng()%80; std::vector<uint8_t> b(len); for(auto&x:b)x=(uint8_t)rng(); uint8_t* d=(uint8_t*)malloc(len?len:1); if(len) memcpy(d,b.data(),len);
        uint64_t nb=rng()%700; std::vector<uint8_t> bits(nb+1); vsp_codec_bitarray_read(d,len,nb,bits.data());
        uint64_t depth=rng()%4; std::vector<uint64_t> dg(4*((2u<<depth)-1)); vsp_codec_merkle_tree_read(d,len,depth,dg.data()); free(d);}
    printf("accepted %ld rejected %ld\n",ok,bad);
    return 0;
}
