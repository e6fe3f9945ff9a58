"""numpy views of the POD layouts in include/clrt_types.h (byte-for-byte the reference's structs)."""
import numpy as np

FLOAT4 = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("w", "<f4")])
UINT2 = np.dtype([("x", "<u4"), ("y", "<u4")])
AABB = np.dtype([("bounds", "<f4", (2, 4))])
BVH_NODE = np.dtype([("data", "<u4", (4,)), ("bounds", "<f4", (2, 4))])                     # 48 B
CONTACT = np.dtype([("pixelIndex", "<u4"), ("materialIndex", "<u4"), ("pad", "<u4", (2,)),
                    ("nd", "<f4", (4,))])                                                    # 32 B
RAY = np.dtype([("idx", "<u4"), ("_pad", "<u4", (3,)), ("origin", "<f4", (4,)), ("direction", "<f4", (4,))])  # 48 B
CAMERA = np.dtype([("FOVDistance", "<f4"), ("resX", "<u4"), ("resY", "<u4"), ("supersampilngFactor", "<u4"),
                   ("viewTransform", "<f4", (12,))])                                         # 64 B
LIGHT = np.dtype([("posAndEnergy", "<f4", (4,))])
MATERIAL = np.dtype([("ambient", "<f4", (4,)), ("diffuse", "<f4", (4,)), ("specular", "<f4", (4,)),
                     ("transmittance", "<f4", (4,)), ("emission", "<f4", (4,)),
                     ("shininess", "<f4"), ("ior", "<f4"), ("dissolve", "<f4"), ("illum", "<f4")])   # 96 B
SCENE_HEADER = np.dtype([("totalDataSize", "<u8"), ("numberOfPrimitives", "<u8"), ("numberOfLights", "<u8"),
                         ("numberOfSpheres", "<u8"), ("numberOfMaterials", "<u8"), ("modelBufferSize", "<u8"),
                         ("numberOfModels", "<u8"), ("totalNumberOfTriangles", "<u8"),
                         ("modelsBoundingBox", "<f4", (2, 4))])                              # 96 B
MODEL_HEADER = np.dtype([("dataSize", "<u8"), ("numberOfSubmeshes", "<u8"), ("numberOfTriangles", "<u8"), ("pad", "<u8"),
                         ("boundingBox", "<f4", (2, 4))])                                    # 64 B
MESH_HEADER = np.dtype([("dataSize", "<u8"), ("numberOfTriangles", "<u8"), ("numberOfVertices", "<u8"),
                        ("numberOfIndices", "<u8"), ("materialIndex", "<u8"), ("pad", "<u8")])   # 48 B
GRID_DATA = np.dtype([("resX", "<u4"), ("resY", "<u4"), ("resZ", "<u4"), ("stepX", "<f4"), ("stepY", "<f4"), ("stepZ", "<f4"),
                      ("leafDensity", "<f4"), ("padTo64", "<f4"), ("box", "<f4", (2, 4))])   # 64 B
TOP_CELL = np.dtype([("resX", "<u4"), ("resY", "<u4"), ("resZ", "<u4"), ("firstLeafIdx", "<u4")])   # 16 B
TRIANGLE = np.dtype([("v0", "<f4", (3,)), ("material", "<u4"), ("v1", "<f4", (3,)), ("submesh", "<u4"),
                     ("v2", "<f4", (3,)), ("model", "<u4")])                                 # 48 B

SIZES = {"BVH_NODE": 48, "AABB": 32, "CONTACT": 32, "RAY": 48, "CAMERA": 64, "LIGHT": 16, "MATERIAL": 96,
         "SCENE_HEADER": 96, "MODEL_HEADER": 64, "MESH_HEADER": 48, "GRID_DATA": 64, "TOP_CELL": 16, "TRIANGLE": 48,
         "UINT2": 8, "FLOAT4": 16}
for _k, _v in SIZES.items():
    assert globals()[_k].itemsize == _v, (_k, globals()[_k].itemsize, _v)
