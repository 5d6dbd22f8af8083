"""Synthetic inputs shared by bench.py and the tests (plain numpy; imports neither the product nor the oracles).

Reference assets are not redistributed, and the cup / backpack meshes the reference's scenes name are not in its tree
(SURVEY §8d C4), so the benchmark frames use seeded stand-ins of the same shape: an earth-map-sized RGB8 image, a
procedural bumpy-sphere triangle mesh with per-vertex normals and uvs, a procedural mesh texture."""
import numpy as np

SEED = 1984     # Philox seed of every bench render; the reference hard-codes the same number for cuRAND (render.h:91)

# BASELINE.json configs: name -> (catalog scene, width, aspect, spp per GPU, description)
WORKLOADS = {
    "final": ("final", 1920, 16.0 / 9.0, 1024, "Next-Week final scene 1920x1079 1024spp/GPU depth50 (BASELINE config 3)"),
    "cornell_smoke": ("cornell_smoke", 800, 1.0, 1024, "Cornell box with smoke 800x800 1024spp/GPU depth50 (BASELINE config 2)"),
    "big1": ("big1", 400, 16.0 / 9.0, 16, "RTIOW random spheres 400x224 16spp depth50 (BASELINE config 1)"),
    "two_perlin": ("two_perlin", 1280, 16.0 / 9.0, 256, "two Perlin-noise spheres 1280x719 256spp/GPU depth50 (scenes.h:248; texture-bound shade)"),
    "mesh": ("mesh", 1920, 16.0 / 9.0, 256, "procedural 1M-triangle mesh + ground 1920x1079 256spp/GPU (BASELINE config 4 STAND-IN: the "
                                           "reference's cup/backpack assets are not in its tree; door_scene's asset is not shipped to the GPU box)"),
    "final_c5": ("final", 3840, 16.0 / 9.0, 4096, "Next-Week final scene 3840x2159 4096spp TOTAL depth50, strong scaling (BASELINE config 5)"),
}


def image_height(width, aspect):
    """render_settings::calc_height (render.h:32-33): int(width / aspect) with the aspect ratio held as a double that was
    assigned from a float (1920 at 16:9 gives 1079)."""
    return int(int(width) / float(np.float32(aspect)))


def synthetic_earth(w=3410, h=1518):
    """Stand-in for textures/earthmap.jpg (same size): smooth continents + noise, RGB8."""
    rng = np.random.default_rng(2026)
    y, x = np.mgrid[0:h, 0:w].astype(np.float32)
    land = np.sin(x * 0.004) * np.cos(y * 0.006) + 0.5 * np.sin(x * 0.013 + y * 0.009)
    img = np.empty((h, w, 3), np.float32)
    img[..., 0] = np.where(land > 0.2, 90 + 60 * land, 20)
    img[..., 1] = np.where(land > 0.2, 120 + 40 * land, 60 + 30 * land)
    img[..., 2] = np.where(land > 0.2, 60, 150 + 40 * land)
    img += rng.integers(-8, 9, size=img.shape)
    return np.clip(img, 0, 255).astype(np.uint8)


def procedural_image(w, h, seed=7):
    """Synthetic RGB8 texture standing in for earthmap.jpg / mesh textures (reference assets are not copied)."""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w].astype(np.float32)
    img = np.stack([127 + 120 * np.sin(x * 0.021) * np.cos(y * 0.017), 127 + 120 * np.sin(x * 0.013 + y * 0.011),
                    127 + 120 * np.cos(x * 0.007 - y * 0.019)], axis=-1)
    img += rng.integers(-6, 7, size=img.shape)
    return np.clip(img, 0, 255).astype(np.uint8)


def procedural_mesh(n_theta, n_phi, seed=3):
    """Seeded bumpy sphere mesh (stands in for the missing cup/backpack .obj): flat arrays as create_meshes_d
    takes them (vertices 3V, indices 3T, uvs 2V, normals 3V — normals deliberately not unit length)."""
    rng = np.random.default_rng(seed)
    th = np.linspace(0.05, np.pi - 0.05, n_theta, dtype=np.float64)
    ph = np.linspace(0, 2 * np.pi, n_phi, endpoint=False, dtype=np.float64)
    T, P = np.meshgrid(th, ph, indexing="ij")
    r = 0.5 + 0.05 * np.sin(5 * T) * np.cos(4 * P) + 0.01 * rng.random(T.shape)
    x, y, z = r * np.sin(T) * np.cos(P), r * np.cos(T) + 0.6, r * np.sin(T) * np.sin(P)
    verts = np.stack([x, y, z], -1).reshape(-1, 3).astype(np.float32)
    norms = (np.stack([np.sin(T) * np.cos(P), np.cos(T), np.sin(T) * np.sin(P)], -1).reshape(-1, 3) *
             (0.5 + rng.random((n_theta * n_phi, 1)))).astype(np.float32)
    uvs = np.stack([P / (2 * np.pi), 1 - T / np.pi], -1).reshape(-1, 2).astype(np.float32)
    i, j = np.meshgrid(np.arange(n_theta - 1), np.arange(n_phi), indexing="ij")
    a, b = i * n_phi + j, i * n_phi + (j + 1) % n_phi
    c, d = a + n_phi, b + n_phi
    idx = np.stack([a, c, b, b, c, d], -1).reshape(-1).astype(np.uint32)     # two triangles per quad, row by row
    return verts, idx, uvs, norms


MESH_CAMERA = np.asarray([0, 1.2, -3, 0, 0.6, 0, 0, 1, 0, 30, 16.0 / 9.0, 0.0, 10.0, 0, 1], np.float32)


def big_mesh(n_theta=708, n_phi=708):
    """~1M-triangle mesh of the config-4 stand-in."""
    return procedural_mesh(n_theta, n_phi, seed=3)
