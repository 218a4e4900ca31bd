"""pine_b200 -- B200-native (sm_100a) backend for Pine's `PathIntegrator(BlueSampler(spp), depth)`.

Layout: csrc/ (CUDA kernels + the C ABI of include/pine_b200.h), lib/ (built .so), scene.py /
integrator.py (host-side mirror of the reference's script interface, ctypes over the C ABI),
scenes.py (the benchmark scene descriptions), multigpu.py (tile/sample sharding + one NCCL reduce),
gltf.py (scene_from_gltf / load_scene: .glb and .gltf meshes, Uber materials, camera),
prl.py (loader for the PRL subset Pine's scene files use; `python -m pine_b200 scene.pine`).
"""
from .scene import (ACES, AABB, Box, Cone, Diffuse, Disk, Emissive, Film, Glass, Glossy, Mesh, Metal,
                    Rect, Scene, Sphere, Subsurface, ThinLenCamera, Triangle, Uber, Uncharted2, mat4)
from .integrator import Backend, MultiBackend, BlueSampler, HaltonSampler, PathIntegrator, SobolSampler, UniformSampler
from ._abi import PineB200Error
from .gltf import GltfError, load_scene, scene_from_gltf
