// r1cs_proving_key.bin -> proof.bin in C++, with nothing of crypto3 on the host: include/vsp_crypto3_marshalling.hpp reads the
// blobs into the shim's types, zk::prove (include/vsp_crypto3_shim.hpp) runs on the GPU, serialize_proof writes the 192 bytes.
// Built and run by tests/test_gpu_shim.py, which compares the bytes with the Python path and with the oracle.
// usage: blob_prove_check <dir with pk.bin primary.bin aux.bin rs.bin>   -> prints the proof as hex
// exit code 3 without a CUDA device (no CPU fallback), 1 on any other failure
#include <cstdio>
#include <fstream>
#include <iterator>
#include <string>

#include "vsp_crypto3_marshalling.hpp"

using namespace nil::crypto3;
typedef b200::marshaling_policy mp;
typedef zk::snark::r1cs_gg_ppzksnark<algebra::curves::bls12<381>> basic_system;

static mp::blob_type load(const std::string &path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("cannot open " + path);
    return mp::blob_type(std::istreambuf_iterator<char>(f), std::istreambuf_iterator<char>());
}

int main(int argc, char **argv) {
    if (argc < 2) return 2;
    const std::string dir = argv[1];
    try {
        const auto pk = mp::deserialize_pk_crs(load(dir + "/pk.bin"));
        const auto primary = mp::deserialize_scalar_vector(load(dir + "/primary.bin"));
        const auto aux = mp::deserialize_scalar_vector(load(dir + "/aux.bin"));
        const auto rs = mp::deserialize_scalar_vector(load(dir + "/rs.bin"));
        if (rs.size() != 2) throw std::runtime_error("rs.bin holds r and s");
        vsp_ctx *probe = nullptr;
        if (vsp_ctx_create(0, nullptr, &probe) != VSP_OK) {
            std::fprintf(stderr, "a CUDA device is required\n");
            return 3;
        }
        vsp_ctx_destroy(probe);
        const auto proof = zk::prove<basic_system>(pk, primary, aux, rs[0], rs[1]);
        for (std::uint8_t b : mp::serialize_proof(proof)) std::printf("%02x", b);
        std::printf("\n");
    } catch (const std::exception &e) {
        std::fprintf(stderr, "FAIL %s\n", e.what());
        return 1;
    }
    return 0;
}
