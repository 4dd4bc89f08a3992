// POD stand-in so the reference's model.h parses without assimp (oracle build only; mesh I/O is out of scope).
#pragma once
struct aiVector3D { float x, y, z; };
struct aiMatrix4x4 { float m[16]; };
struct aiScene; struct aiNode;
