/* oracle/_ref shim: assimp is not installed; the reference's loader functions only need to
 * compile (the harness never calls them — meshes enter through create_meshes_d's flat arrays). */
#pragma once
#include <algorithm>
#include <string>
struct aiVector3D { float x, y, z; };
struct aiFace { unsigned mNumIndices; unsigned* mIndices; };
struct aiString { const char* C_Str() const { return ""; } };
enum aiTextureType { aiTextureType_DIFFUSE };
struct aiMaterial { unsigned GetTextureCount(aiTextureType) const { return 0; } void GetTexture(aiTextureType, unsigned, aiString*) const {} };
struct aiMesh { unsigned mNumVertices, mNumFaces; aiVector3D* mVertices; aiVector3D* mNormals; aiVector3D* mTextureCoords[8]; aiFace* mFaces; unsigned mMaterialIndex; };
struct aiNode { unsigned mNumMeshes; unsigned* mMeshes; unsigned mNumChildren; aiNode** mChildren; };
struct aiScene { unsigned mFlags; aiNode* mRootNode; aiMesh** mMeshes; aiMaterial** mMaterials; };
enum { aiProcess_Triangulate = 1, aiProcess_GenNormals = 2, AI_SCENE_FLAGS_INCOMPLETE = 1 };
namespace Assimp { struct Importer { const aiScene* ReadFile(const std::string&, unsigned) { return nullptr; } const char* GetErrorString() const { return "assimp unavailable"; } }; }
using std::reverse;
