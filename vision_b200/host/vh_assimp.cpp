// `model` shape import through the reference's own importer library.
//   AssimpParser::load_scene   reference src/importers/assimp_parser.cpp:309-341  (flag set, RVC property)
//   AssimpParser::parse_meshes reference src/importers/assimp_parser.cpp:231-307  (vertex / triangle extraction, quads split 0-1-2 / 0-2-3,
//                                                                                   invalid normals / uvs -> 0)
// Compiled against the vendored assimp built by tools/build_assimp.py (-DVH_WITH_ASSIMP, vision_b200/host/Makefile); without it the
// `model` plugin falls back to the in-tree OBJ loader (file-order primitive ids) and says so in the scene log.
#include "vh_scene.h"
#include <cmath>

#ifdef VH_WITH_ASSIMP
#include <assimp/Importer.hpp>
#include <assimp/postprocess.h>
#include <assimp/scene.h>
#endif

namespace vh {

#ifdef VH_WITH_ASSIMP
bool assimp_available() { return true; }

std::vector<MeshData> load_model_assimp(const std::string &path, bool swap_handed, bool smooth, bool flip_uv, uint32_t subdiv_level) {
    if (subdiv_level != 0) throw std::runtime_error("model " + path + ": subdiv_level != 0 is not built (Assimp::Subdivider)");
    Assimp::Importer importer;
    importer.SetPropertyInteger(AI_CONFIG_PP_RVC_FLAGS, aiComponent_COLORS | aiComponent_BONEWEIGHTS | aiComponent_ANIMATIONS | aiComponent_LIGHTS |
                                                            aiComponent_CAMERAS | aiComponent_TEXTURES | aiComponent_MATERIALS);
    unsigned steps = aiProcess_JoinIdenticalVertices | (smooth ? aiProcess_GenSmoothNormals : aiProcess_GenNormals) | aiProcess_PreTransformVertices |
                     aiProcess_ImproveCacheLocality | aiProcess_FixInfacingNormals | aiProcess_FindInvalidData | aiProcess_FindDegenerates |
                     aiProcess_GenUVCoords | aiProcess_TransformUVCoords | aiProcess_OptimizeMeshes | aiProcess_Triangulate |
                     (flip_uv ? (unsigned)aiProcess_FlipUVs : 0u);
    if (swap_handed) steps |= aiProcess_MakeLeftHanded | aiProcess_FlipWindingOrder;
    const aiScene *scene = importer.ReadFile(path.c_str(), steps);
    if (!scene) throw std::runtime_error("model " + path + ": " + importer.GetErrorString());
    auto bad = [](float x) { return std::isnan(x) || std::isinf(x); };
    std::vector<MeshData> out;
    out.reserve(scene->mNumMeshes);
    for (unsigned mi = 0; mi < scene->mNumMeshes; ++mi) {
        const aiMesh *am = scene->mMeshes[mi];
        MeshData m;
        m.pos.reserve(3 * (size_t)am->mNumVertices); m.nrm.reserve(3 * (size_t)am->mNumVertices); m.uv.reserve(2 * (size_t)am->mNumVertices);
        for (unsigned i = 0; i < am->mNumVertices; ++i) {
            const aiVector3D &p = am->mVertices[i];
            m.pos.insert(m.pos.end(), {p.x, p.y, p.z});
            float n[3] = {0.f, 0.f, 0.f};
            if (am->mNormals) { n[0] = am->mNormals[i].x; n[1] = am->mNormals[i].y; n[2] = am->mNormals[i].z; }
            if (bad(n[0]) || bad(n[1]) || bad(n[2])) n[0] = n[1] = n[2] = 0.f;
            m.nrm.insert(m.nrm.end(), {n[0], n[1], n[2]});
            float t[2] = {0.f, 0.f};
            if (am->mTextureCoords[0]) { t[0] = am->mTextureCoords[0][i].x; t[1] = am->mTextureCoords[0][i].y; }
            if (bad(t[0]) || bad(t[1])) t[0] = t[1] = 0.f;
            m.uv.insert(m.uv.end(), {t[0], t[1]});
        }
        for (unsigned f = 0; f < am->mNumFaces; ++f) {
            const aiFace &face = am->mFaces[f];
            if (face.mNumIndices == 3) m.tri.insert(m.tri.end(), {face.mIndices[0], face.mIndices[1], face.mIndices[2]});
            else if (face.mNumIndices == 4) {
                m.tri.insert(m.tri.end(), {face.mIndices[0], face.mIndices[1], face.mIndices[2]});
                m.tri.insert(m.tri.end(), {face.mIndices[0], face.mIndices[2], face.mIndices[3]});
            }       // points / lines left by FindDegenerates are skipped, as in the reference
        }
        out.push_back(std::move(m));
    }
    return out;
}
#else
bool assimp_available() { return false; }
std::vector<MeshData> load_model_assimp(const std::string &path, bool, bool, bool, uint32_t) {
    throw std::runtime_error("model " + path + ": libvision_host was built without the vendored assimp (tools/build_assimp.py)");
}
#endif

}  // namespace vh
