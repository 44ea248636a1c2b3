"""The per-frame body of the reference's fusion loop (apps/fusion/pipeline.py:322-414) with the dense-depth fitter in the place of the
neural tracker (`run_non_rigid_alignment`, pipeline.py:356-367) -- SURVEY.md §8(f) rank 3 "pipeline wiring". Headless and on the device:
every step is a call into the `nnrt` tree (either front end), no tensor leaves the GPU between integrate -> extract -> fit.

    frame 0   blocks touched by the depth surface  -> VoxelBlockGrid.integrate (rigid)            pipeline.py: volume.integrate(...)
              canonical mesh                        -> extract_triangle_mesh
              graph nodes                           -> median-grid subsample of the mesh vertices -> HierarchicalGraphWarpField
    frame k   blocks the moved surface will reach   -> find_blocks_intersecting_truncation_region  pipeline.py:327-334
              non-rigid alignment                   -> DeformableMeshToImageFitter.fit_to_image    (replaces pipeline.py:356-367)
              fuse                                  -> integrate_non_rigid through the fitted field pipeline.py:395-400
              canonical + warped mesh               -> extract_triangle_mesh, warp_mesh            pipeline.py:408-409

Image conventions. The TSDF side (Open3D's) projects with the pinhole model, pixel centres at integers, +y down: u = fx x/z + cx,
v = fy y/z + cy. The fitter's rasteriser (the reference's, kept bit for bit: SURVEY.md "y-flip in NDC, cx sign") puts pixel centres at
half integers and +y UP: u = fx x/z - cx + W - 1/2, v = -fy y/z + cy - 1/2. A mesh extracted from the volume and handed to the fitter
together with the depth frame it came from would therefore be drawn upside down against it (the reference never wires the two
together; its fitter tests use symmetric planes). The loop reconciles the two without touching either side: the fitter works in the
MIRRORED frame y -> -y (mesh and nodes reflected, triangle winding swapped so front faces stay front) with the principal point
(W - 1/2 - cx, cy + 1/2), under which both projections agree pixel for pixel; the fitted node motions are mapped back (R = S R' S,
t = S t', S = diag(1, -1, 1)) into the volume's frame for integration and warping.

Left out, as in DESIGN.md §7: Open3D's RGB-D odometry (rigid alignment; the extrinsics are an input here), DeformNet, telemetry, keyframe
policy, cropping. The reference's non-rigid integration never raises a voxel's weight (NonRigidSurfaceVoxelBlockGridImpl.h:201-203), so
only blocks of the first (rigidly integrated) frame ever pass a positive weight threshold; `weight_threshold` defaults to 0.5 for that reason."""
import numpy as np
import torch

from . import nnrt as _default_nnrt


class FusionLoop:
    def __init__(self, intrinsics, extrinsics=None, voxel_size=0.004, block_resolution=8, block_count=4096, node_spacing=0.03, node_coverage=None,
                 anchor_count=4, layer_count=2, depth_scale=1000.0, far_clipping_distance=3.0, truncation_voxel_multiplier=2.0, weight_threshold=0.5,
                 fitter=None, nnrt=None, warm_start_rotations=False):
        """Solver settings: with the reference's defaults (arap_term_weight 200, no damping) and overlapping anchor weights the reference's
        iteration diverges on multi-node graphs (dynamicfuion_python_b200/synthetic.py, DESIGN.md "reference behaviour"); the defaults here
        are the contractive ones the parity scenes use: node coverage a quarter of the node spacing, arap_term_weight 20, lambda 1."""
        self.nnrt = _default_nnrt if nnrt is None else nnrt
        G = self.nnrt.geometry
        self.K = np.asarray(intrinsics, np.float64).reshape(3, 3)
        self.extrinsics = np.eye(4) if extrinsics is None else np.asarray(extrinsics, np.float64).reshape(4, 4)
        self.volume = G.NonRigidSurfaceVoxelBlockGrid(["tsdf", "weight", "color"], [torch.float32, torch.float32, torch.float32], [(1,), (1,), (3,)],
                                                      voxel_size, block_resolution, block_count)
        node_coverage = node_spacing / 4.0 if node_coverage is None else node_coverage
        self.node_spacing, self.node_coverage, self.anchor_count, self.layer_count = node_spacing, node_coverage, anchor_count, layer_count
        self.depth_scale, self.far, self.truncation, self.weight_threshold = depth_scale, far_clipping_distance, truncation_voxel_multiplier, weight_threshold
        self.fitter = fitter  # None: made in first_frame, when the image size is known (the ARAP weight follows it)
        # Each frame is fitted canonical -> current, so the node state of the previous frame is only a starting point. Translations are
        # carried over; rotations restart from the identity by default: the reference iteration is not contractive in warm-started
        # rotations (bench.py makes the same choice; measured on the 640 x 480 relief: divergence within three frames otherwise).
        self.warm_start_rotations = warm_start_rotations
        self.graph = self._fit_graph = self.canonical_mesh = self.warped_mesh = None
        self.frame_index = 0

    def _depth_normals(self, depth):
        """Normal map of the depth frame, oriented AWAY from the camera: the reference's kernel integrates a voxel only where
        dot(direction to the camera, normal) <= 0.5 (NonRigidSurfaceVoxelBlockGridImpl.h:192)."""
        f = self.nnrt.geometry.functional
        d = torch.as_tensor(np.asarray(depth)) if not isinstance(depth, torch.Tensor) else depth
        points, _ = f.unproject_raster_depth_without_filtering(d, self.K, None, self.depth_scale, self.far, True)
        h, w = points.shape[0], points.shape[1]
        normals = f.compute_ordered_point_cloud_normals(points.reshape(-1, 3), (h, w)).reshape(h, w, 3)
        towards_camera = (normals * points).sum(-1, keepdim=True) < 0
        return torch.where(towards_camera, -normals, normals).contiguous()

    def first_frame(self, depth, color=None):
        """Rigid integration of frame 0, canonical mesh, graph: returns the canonical mesh."""
        G = self.nnrt.geometry
        blocks = self.volume.compute_unique_block_coordinates(depth, self.K, self.extrinsics, self.depth_scale, self.far, self.truncation)
        if color is None:
            self.volume.integrate(blocks, depth, self.K, self.extrinsics, self.depth_scale, self.far, self.truncation)
        else:
            self.volume.integrate(blocks, depth, color, self.K, self.K, self.extrinsics, self.depth_scale, self.far, self.truncation)
        self.canonical_mesh = self.volume.extract_triangle_mesh(self.weight_threshold)
        vertices = self.canonical_mesh.positions
        nodes = vertices[G.functional.median_grid_subsample_3d_points(vertices, self.node_spacing)].contiguous()
        # coarser layers are decimated relative to the node SPACING (the class default scales the node coverage, which here is a
        # quarter of the spacing and would leave every layer the size of the first)
        spacing = self.node_spacing
        self.graph = G.HierarchicalGraphWarpField(nodes, self.node_coverage, False, self.anchor_count, 0,
                                                  G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, self.layer_count, 4,
                                                  lambda i_layer, coverage: spacing * 2.0 ** i_layer)
        # the same hierarchy in the fitter's mirrored frame (see "Image conventions")
        self._mirror = torch.tensor([1.0, -1.0, 1.0], device=nodes.device)
        self._mirror_block = self._mirror[:, None] * self._mirror[None, :]
        g = self.graph
        self._fit_graph = G.HierarchicalGraphWarpField.from_graph(
            (g.get_node_positions(True) * self._mirror).contiguous(), g.get_edges(), g.first_layer_node_count, self.node_coverage, False, self.anchor_count, 0,
            G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, g.get_edge_layer_indices(), g.get_layer_decimation_radii())
        height, width = (int(x) for x in (depth.shape[0], depth.shape[1]))
        if self.fitter is None:
            # the data term grows with the number of pixels a node sees, the ARAP term does not: the weight that keeps the iteration
            # contractive at 160 x 120 (20) is scaled by the square root of the pixel ratio (80 at 640 x 480; measured stable over 8
            # frames there with 20 mm node spacing, divergent at 20 and at 200 -- tests/measure/fusion_loop_times.py)
            self.fitter = self.nnrt.alignment.DeformableMeshToImageFitter(
                max_iteration_count=10, preconditioning_dampening_factor=1.0, arap_term_weight=20.0 * float(np.sqrt(height * width / 19200.0)))
        self._fit_K = self.K.copy()
        self._fit_K[0, 2], self._fit_K[1, 2] = width - 0.5 - self.K[0, 2], self.K[1, 2] + 0.5
        self.warped_mesh, self.frame_index = self.canonical_mesh, 1
        return self.canonical_mesh

    def _mirrored_mesh(self, mesh):
        return ((mesh.positions * self._mirror).contiguous(), (mesh.normals * self._mirror).contiguous(), mesh.triangles[:, [0, 2, 1]].contiguous())

    def next_frame(self, depth, color=None):
        """One tracked and fused frame: returns (canonical_mesh, warped_mesh, cos_voxel_ray_to_normal)."""
        if self.graph is None:
            raise RuntimeError("FusionLoop.next_frame: call first_frame first")
        if not np.allclose(self.extrinsics, np.eye(4)):
            raise RuntimeError("FusionLoop: the tracker runs in the camera frame; non-identity extrinsics are not wired (rigid odometry is out of scope)")
        blocks = self.volume.find_blocks_intersecting_truncation_region(depth, self.graph, self.K, self.extrinsics, self.depth_scale, self.far,
                                                                        self.truncation)
        self.volume.hashmap().activate(blocks)
        # non-rigid alignment in the mirrored frame, warm-started from the field's current state; result mapped back
        fit = self._fit_graph
        if self.warm_start_rotations:
            fit.set_node_rotations((self.graph.get_node_rotations(True) * self._mirror_block).contiguous(), True)
        else:
            fit.reset_rotations()
        fit.set_node_translations((self.graph.get_node_translations(True) * self._mirror).contiguous(), True)
        self.fitter.fit_to_image(fit, self._mirrored_mesh(self.canonical_mesh), color, depth, None, self._fit_K, None, self.depth_scale)
        self.fitter.check_status()
        self.graph.set_node_rotations((fit.get_node_rotations(True) * self._mirror_block).contiguous(), True)
        self.graph.set_node_translations((fit.get_node_translations(True) * self._mirror).contiguous(), True)
        empty_color = torch.zeros((0, 0, 3), dtype=torch.float32)
        cosines = self.volume.integrate_non_rigid(blocks, self.graph, depth, empty_color if color is None else color, self._depth_normals(depth),
                                                  self.K, self.K, self.extrinsics, self.depth_scale, self.far, self.truncation)
        self.canonical_mesh = self.volume.extract_triangle_mesh(self.weight_threshold)
        self.warped_mesh = self.graph.warp_mesh(self.canonical_mesh)
        self.frame_index += 1
        return self.canonical_mesh, self.warped_mesh, cosines
