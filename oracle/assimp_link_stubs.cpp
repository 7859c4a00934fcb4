/*
 * Link-time stand-ins for the four Assimp::Importer symbols MeshEngine references
 * (meshEngine.h:56, meshEngine.cpp:141-145). TEST INFRASTRUCTURE ONLY.
 * The Assimp library itself is not built here (it needs its cmake-generated
 * headers); the harness injects scenes as hand-filled aiScene structs instead,
 * so ReadFile() is never the source of geometry and simply fails.
 */
#include "assimp/Importer.hpp"
namespace Assimp {
Importer::Importer() : pimpl(nullptr) {}
Importer::~Importer() {}
const aiScene* Importer::ReadFile(const char*, unsigned int) { return nullptr; }
const char* Importer::GetErrorString() const { return "Assimp is not linked into oracle/_ref (scene injection only)"; }
} // namespace Assimp
