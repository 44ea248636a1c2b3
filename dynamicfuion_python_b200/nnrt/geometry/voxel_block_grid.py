"""nnrt.geometry.VoxelBlockGrid / NonRigidSurfaceVoxelBlockGrid (cpp/pybind/geometry/geometry.cpp:61-276,
cpp/geometry/NonRigidSurfaceVoxelBlockGrid.cpp:28-260, cpp/geometry/VoxelBlockGrid.cpp): the TSDF volume on the other side of the
tracker in the fusion loop -- integrate the depth frame through the warp field, extract the canonical mesh, fit.

The reference keeps the blocks in Open3D's device hash map, which is absent here. The volume is instead
  * a key buffer [capacity, 3] int32 and one value buffer per attribute [capacity, R, R, R, channels], all on the device;
  * blocks are only ever appended (a TSDF volume never deactivates a block), so a block's buffer index is its rank of activation and
    the active indices are 0 .. size-1;
  * membership queries go through the C ABI's block hash (nnrt_b200_voxel_block_neighbors: open addressing over packed keys, built per
    call), which also gives the 27-neighbourhood that block activation and mesh extraction need.
Every per-voxel / per-block computation is a CUDA kernel behind include/nnrt_b200.h (integration, warped-block boxes, surface mask, block
hash, marching cubes); torch only holds the buffers and does the boolean indexing between the calls. There is no CPU path."""
import torch

from .. import _tensor as T
from . import functional

_DTYPE_NAMES = {"Float32": torch.float32, "Float64": torch.float64, "UInt16": torch.uint16, "Int16": torch.int16, "UInt8": torch.uint8,
                "Int32": torch.int32, "Int64": torch.int64, "Bool": torch.bool}


def _dtype(value):
    if isinstance(value, torch.dtype):
        return value
    name = str(value).split(".")[-1]
    if name in _DTYPE_NAMES:
        return _DTYPE_NAMES[name]
    raise RuntimeError(f"unsupported voxel attribute dtype {value!r}")


def _raw(tensor):
    """uint16 tensors as int16 for the tensor ops torch does not implement for unsigned 16-bit (indexing, concatenation); same bits."""
    return tensor.view(torch.int16) if tensor.dtype == torch.uint16 else tensor


class BlockHashMap:
    """What the grid classes and nnrt.io use of open3d.core.HashMap: key_tensor() [capacity,3] int32, value_tensor(i), active_buf_indices()
    int32 [size], size(), capacity(), activate(keys) / insert(keys, values) / find(keys) returning (buffer indices int32, masks bool)."""

    def __init__(self, capacity, value_dtypes, value_element_shapes, device):
        self._capacity, self._size, self._device = max(int(capacity), 1), 0, device
        self._keys = torch.zeros((self._capacity, 3), dtype=torch.int32, device=device)
        self._values = [torch.zeros((self._capacity,) + tuple(shape), dtype=dtype, device=device) for dtype, shape in zip(value_dtypes, value_element_shapes)]

    def size(self):
        return self._size

    def capacity(self):
        return self._capacity

    def get_device(self):
        return self._device

    def key_tensor(self):
        return self._keys

    def value_tensor(self, index=0):
        return self._values[index]

    def value_tensors(self):
        return list(self._values)

    def active_buf_indices(self):
        return torch.arange(self._size, dtype=torch.int32, device=self._device)

    def _reserve(self, count):
        if count <= self._capacity:
            return
        capacity = max(count, 2 * self._capacity)  # Open3D rehashes to a larger table; buffer indices of existing blocks are kept here
        grow = capacity - self._capacity
        self._keys = torch.cat([self._keys, torch.zeros((grow, 3), dtype=torch.int32, device=self._device)])
        self._values = [torch.cat([_raw(v), torch.zeros((grow,) + tuple(v.shape[1:]), dtype=_raw(v).dtype, device=self._device)]).view(v.dtype)
                        for v in self._values]
        self._capacity = capacity

    def find(self, keys):
        """Buffer index of each key [M,3] (int32; undefined where the mask is False) and the found mask."""
        keys = T.cuda(keys, torch.int32).reshape(-1, 3)
        if keys.shape[0] == 0 or self._size == 0:
            return torch.zeros((keys.shape[0],), dtype=torch.int32, device=self._device), torch.zeros((keys.shape[0],), dtype=torch.bool, device=self._device)
        # one hash over [active keys; queries]: slot 13 (offset 0,0,0) of a query row is the first row with that key
        first = functional.voxel_block_neighbors(torch.cat([self._keys[:self._size], keys]))[self._size:, 13]
        return first.clamp(max=self._size - 1), first < self._size

    def activate(self, keys):
        """Appends the keys that are not yet in the map (each once, in order of first appearance); returns (buffer indices, masks) like
        HashMap::Activate: masks[i] is True where key i was newly activated by this call (its first appearance)."""
        keys = T.cuda(keys, torch.int32).reshape(-1, 3)
        count = keys.shape[0]
        if count == 0:
            return torch.zeros((0,), dtype=torch.int32, device=self._device), torch.zeros((0,), dtype=torch.bool, device=self._device)
        first = functional.voxel_block_neighbors(torch.cat([self._keys[:self._size], keys]))[self._size:, 13]
        rows = torch.arange(self._size, self._size + count, dtype=torch.int32, device=self._device)
        fresh = first == rows  # not among the active keys and the first of its kind in this batch
        new_count = int(fresh.sum())
        self._reserve(self._size + new_count)
        position = torch.cumsum(fresh.to(torch.int32), 0, dtype=torch.int32) - 1 + self._size  # buffer index of each fresh key
        self._keys[self._size:self._size + new_count] = keys[fresh]
        # every key's buffer index: an active block's own, else that of the fresh key it repeats
        repeated = (first.long() - self._size).clamp(min=0)
        indices = torch.where(first < self._size, first, position[repeated])
        self._size += new_count
        return indices, fresh

    def insert(self, keys, values):
        """Activates `keys` and stores values[i] rows for the newly activated ones (HashMap::Insert keeps the first value of a key)."""
        indices, fresh = self.activate(keys)
        target = indices[fresh].long()
        for buffer, value in zip(self._values, values):
            rows = _raw(T.to_torch(value).to(self._device))[fresh]
            _raw(buffer)[target] = rows.to(_raw(buffer).dtype).reshape((-1,) + tuple(buffer.shape[1:]))
        return indices, fresh


class VoxelBlockGrid:
    """A sparse grid of dense voxel blocks (cpp/geometry/VoxelBlockGrid.h). attr_names / attr_dtypes / attr_channels as in the reference
    (e.g. ("tsdf", "weight", "color"), (Float32, UInt16, Float32), ((1,), (1,), (3,)))."""

    def __init__(self, attr_names=(), attr_dtypes=(), attr_channels=(), voxel_size=0.0058, block_resolution=16, block_count=10000, device=None):
        if not (len(attr_names) == len(attr_dtypes) == len(attr_channels)):
            raise RuntimeError("Number of attribute names, dtypes, and channels mismatch.")  # VoxelBlockGrid.cpp
        if float(voxel_size) <= 0 or int(block_resolution) <= 0:
            raise RuntimeError("voxel_size and block_resolution must be positive.")
        self.voxel_size, self.block_resolution = float(voxel_size), int(block_resolution)
        self._names = list(attr_names)
        self._device = T.device() if device is None else torch.device(device)
        if self._device.type != "cuda":
            raise RuntimeError("nnrt (B200 build): voxel block grids live on a CUDA device; there is no CPU path")
        r = self.block_resolution
        shapes = []
        for channels in attr_channels:
            channels = (int(channels),) if isinstance(channels, int) else tuple(int(c) for c in channels)
            shapes.append((r, r, r) + channels)
        self._hashmap = BlockHashMap(block_count, [_dtype(d) for d in attr_dtypes], shapes, self._device)

    def _assert_initialized(self):
        if not self._names:
            raise RuntimeError("VoxelBlockGrid not initialized.")

    def hashmap(self):
        return self._hashmap

    def attribute_names(self):
        return list(self._names)

    def attribute(self, attr_name):
        self._assert_initialized()
        if attr_name not in self._names:
            raise RuntimeError(f"Attribute {attr_name} not found.")
        return self._hashmap.value_tensor(self._names.index(attr_name))

    def get_device(self):
        return self._device

    def get_block_resolution(self):
        return self.block_resolution

    def get_block_count(self):
        return self._hashmap.capacity()

    def get_voxel_size(self):
        return self.voxel_size

    def _value_views(self):
        """The dict the kernels take: tsdf / weight as [capacity,R,R,R], colour as [capacity,R,R,R,3] (views, updated in place)."""
        r = self.block_resolution
        views = {}
        for name in ("tsdf", "weight", "color"):
            if name in self._names:
                buffer = self.attribute(name)
                views[name] = buffer.view(buffer.shape[0], r, r, r, 3) if name == "color" else buffer.view(buffer.shape[0], r, r, r)
        if "tsdf" not in views or "weight" not in views:
            raise RuntimeError("the TSDF operations need 'tsdf' and 'weight' attributes")
        return views

    def compute_unique_block_coordinates(self, depth, intrinsic, extrinsic=None, depth_scale=1000.0, depth_max=3.0, trunc_voxel_multiplier=8.0):
        """Blocks touched by the truncation band around the depth surface: Open3D's GetUniqueBlockCoordinates (re-exported at
        pybind/geometry/geometry.cpp:152-168; third party, restated from the published DepthTouch kernel: every 4th pixel in each
        direction, four samples along the pixel's ray from depth - truncation to depth + truncation, block = floor(point / block
        size)). Returns [M,3] int32, each block once, in order of first appearance (Open3D's order is that of its hash map). Frame
        bootstrap, not a per-iteration path: tensor ops + one pass through the block hash."""
        self._assert_initialized()
        d = T.to_torch(depth.as_tensor() if hasattr(depth, "as_tensor") else depth)
        if d.dim() == 3:
            d = d[..., 0]
        if d.dtype == torch.uint16:
            d = d.view(torch.int16).to(torch.int32) & 0xFFFF
        d = d.to(self._device).to(torch.float32)[::4, ::4] / float(depth_scale)
        K = T.host_f64(intrinsic, (3, 3))
        pose = torch.linalg.inv(T.host_f64(torch.eye(4, dtype=torch.float64) if extrinsic is None else extrinsic, (4, 4)))  # camera -> world
        rows, cols = d.shape
        v, u = torch.meshgrid(torch.arange(rows, device=self._device, dtype=torch.float32) * 4, torch.arange(cols, device=self._device, dtype=torch.float32) * 4,
                              indexing="ij")
        valid = (d > 0) & (d < float(depth_max))
        rays = torch.stack([(u - float(K[0, 2])) / float(K[0, 0]), (v - float(K[1, 2])) / float(K[1, 1]), torch.ones_like(u)], -1)[valid]  # z = 1
        rotation, origin = pose[:3, :3].to(self._device, torch.float32), pose[:3, 3].to(self._device, torch.float32)
        directions = rays @ rotation.t()
        depth_valid = d[valid]
        truncation = self.voxel_size * float(trunc_voxel_multiplier)
        t_min, t_max = (depth_valid - truncation).clamp(min=0.0), (depth_valid + truncation).clamp(max=float(depth_max))
        steps = torch.arange(4, device=self._device, dtype=torch.float32)
        t = t_min[:, None] + steps[None, :] * ((t_max - t_min) / 3.0)[:, None]
        points = origin[None, None, :] + t[..., None] * directions[:, None, :]
        blocks = torch.floor(points / (self.voxel_size * self.block_resolution)).to(torch.int32).reshape(-1, 3)
        scratch = BlockHashMap(max(int(blocks.shape[0]), 1), [], [], self._device)
        scratch.activate(blocks)
        return scratch.key_tensor()[:scratch.size()].clone()

    def integrate(self, block_coords, depth, *args):
        """integrate(block_coords, depth, color, depth_intrinsic, color_intrinsic, extrinsic, depth_scale=1000, depth_max=3,
        trunc_voxel_multiplier=8) or integrate(block_coords, depth, intrinsic, extrinsic, depth_scale=1000, depth_max=3,
        trunc_voxel_multiplier=8) -- the overloads the reference re-exports from Open3D (pybind/geometry/geometry.cpp:170-203): activates
        block_coords and integrates the frame rigidly into THOSE blocks (csrc/tsdf.cu::integrate_rigid_kernel)."""
        self._assert_initialized()
        args = list(args)
        first = T.to_torch(args[0].as_tensor() if hasattr(args[0], "as_tensor") else args[0]) if args and args[0] is not None else None
        if first is not None and tuple(first.shape) == (3, 3):  # depth-only overload
            color, depth_intrinsic, color_intrinsic, rest = None, args[0], args[0], args[1:]
        else:
            color, depth_intrinsic, color_intrinsic, rest = args[0], args[1], args[2], args[3:]
        extrinsic = rest[0]
        depth_scale, depth_max, multiplier = (list(rest[1:]) + [1000.0, 3.0, 8.0][len(rest) - 1:])[:3]
        indices, _ = self._hashmap.activate(block_coords)
        indices = torch.unique(indices)  # a block listed twice is integrated once
        functional.integrate_voxel_blocks(indices.long(), self._hashmap.key_tensor(), self._value_views(), self.block_resolution, self.voxel_size, depth,
                                          color, depth_intrinsic, color_intrinsic, extrinsic, depth_scale, depth_max, multiplier)

    def extract_triangle_mesh(self, weight_threshold=3.0, estimated_vertex_number=-1):
        """Specific operation for TSDF volumes. Extract triangle mesh at isosurface points (VoxelBlockGrid.cpp:461-484); marching cubes
        on the device over the active blocks, neighbour blocks through the block hash. `estimated_vertex_number` is accepted and unused:
        the vertex count is found by a device scan."""
        self._assert_initialized()
        size = self._hashmap.size()
        views = {name: value[:size] for name, value in self._value_views().items()}
        return functional.extract_triangle_mesh(self._hashmap.key_tensor()[:size], views, self.voxel_size, weight_threshold)


class NonRigidSurfaceVoxelBlockGrid(VoxelBlockGrid):
    """cpp/geometry/NonRigidSurfaceVoxelBlockGrid.h: a TSDF volume in the canonical frame, integrated through a warp field."""

    def buffer_coordinates_of_inactive_neighbor_blocks(self):
        """BufferCoordinatesOfInactiveNeighborBlocks (.cpp:76-106): for every active block and each of its 27 neighbour offsets, in
        (offset, block) order, the coordinates that are not active -- with repeats, as the reference returns them."""
        size = self._hashmap.size()
        keys = self._hashmap.key_tensor()[:size]
        if size == 0:
            return torch.zeros((0, 3), dtype=torch.int32, device=self._device)
        table = functional.voxel_block_neighbors(keys)  # [B,27], slot (dz + 1) * 9 + (dy + 1) * 3 + (dx + 1) as the reference's neighbor_index
        slots = torch.arange(27, device=self._device)
        offsets = torch.stack([slots % 3 - 1, (slots % 9) // 3 - 1, slots // 9 - 1], 1).to(torch.int32)  # (dx, dy, dz)
        candidates = (keys[None, :, :] + offsets[:, None, :]).reshape(-1, 3)
        return candidates[(table.t() < 0).reshape(-1)]

    def activate_sleeve_blocks(self):
        """ActivateSleeveBlocks (.cpp:108-118): activates the inactive neighbours of the active blocks; returns how many coordinates
        were offered (repeats included, as the reference counts them)."""
        coordinates = self.buffer_coordinates_of_inactive_neighbor_blocks()
        self._hashmap.activate(coordinates)
        return int(coordinates.shape[0])

    def get_bounding_boxes_of_warped_blocks(self, block_keys, warp_field, extrinsics=None):
        return functional.bounding_boxes_of_warped_blocks(block_keys, warp_field, self.voxel_size, self.block_resolution, extrinsics)

    def get_axis_aligned_boxes_intersecting_surface_mask(self, boxes, depth, intrinsics, depth_scale, depth_max, downsampling_factor=4,
                                                         trunc_voxel_multiplier=8.0):
        return functional.axis_aligned_boxes_intersecting_surface_mask(boxes, depth, intrinsics, depth_scale, depth_max, downsampling_factor,
                                                                       self.voxel_size * trunc_voxel_multiplier)

    def find_blocks_intersecting_truncation_region(self, depth, warp_field, intrinsics, extrinsics, depth_scale, depth_max, truncation_voxel_multiplier):
        """FindBlocksIntersectingTruncationRegion (.cpp:150-178): the inactive neighbours of active blocks whose warped bounding box
        meets the truncation region of the depth surface. As in the reference, the mask is evaluated at every 4th pixel with the
        DEFAULT truncation multiplier of GetAxisAlignedBoxesIntersectingSurfaceMask (8.0): `truncation_voxel_multiplier` is accepted
        and not forwarded (.cpp:171-173)."""
        self._assert_initialized()
        coordinates = self.buffer_coordinates_of_inactive_neighbor_blocks()
        if coordinates.shape[0] == 0:
            return coordinates
        boxes = self.get_bounding_boxes_of_warped_blocks(coordinates, warp_field, extrinsics)
        mask = self.get_axis_aligned_boxes_intersecting_surface_mask(boxes, depth, intrinsics, depth_scale, depth_max, 4)
        return coordinates[mask]

    def integrate_non_rigid(self, block_coords, warp_field, depth, color, depth_normals, depth_intrinsics, color_intrinsics, extrinsics,
                            depth_scale, depth_max, truncation_voxel_multiplier):
        """IntegrateNonRigid (.cpp:32-74): activates `block_coords`, then integrates the frame into EVERY active block through the warp
        field. Returns cos_voxel_ray_to_normal [H,W]."""
        self._assert_initialized()
        coordinates = T.cuda(block_coords, torch.int32)
        if coordinates.dim() != 2 or coordinates.shape[1] != 3:
            raise RuntimeError(f"block_coords must have shape [N, 3], got {tuple(coordinates.shape)}")
        self._hashmap.activate(coordinates)
        active = self._hashmap.active_buf_indices().long()
        return functional.integrate_non_rigid_voxel_blocks(active, self._hashmap.key_tensor(), self._value_views(), self.block_resolution, self.voxel_size,
                                                           warp_field, depth, color, depth_normals, depth_intrinsics, color_intrinsics, extrinsics,
                                                           depth_scale, depth_max, truncation_voxel_multiplier)

    def extract_voxel_block_coordinates(self):
        """ExtractVoxelBlockCoordinates (.cpp:196-201): metric origin of every active block, [B,3] f32."""
        size = self._hashmap.size()
        return self._hashmap.key_tensor()[:size].to(torch.float32) * self.block_resolution * self.voxel_size

    def extract_voxel_values_and_coordinates(self):
        """ExtractVoxelValuesAndCoordinates (.cpp:180-194; kernel/NonRigidSurfaceVoxelBlockGridImpl.h:438-548): per voxel of every active
        block (x, y, z metric, tsdf[, weight][, r, g, b]) as float32, voxels in (block, z, y, x) order. Telemetry: gathered with tensor ops."""
        size, r = self._hashmap.size(), self.block_resolution
        keys = self._hashmap.key_tensor()[:size]
        local = torch.arange(r, device=self._device)
        z, y, x = torch.meshgrid(local, local, local, indexing="ij")
        lattice = torch.stack([x, y, z], -1).reshape(1, -1, 3)
        coordinates = ((keys.long() * r)[:, None, :] + lattice).to(torch.float32) * self.voxel_size
        columns = [coordinates.reshape(-1, 3), self.attribute("tsdf")[:size].reshape(-1, 1).to(torch.float32)]
        if "weight" in self._names:
            weight = self.attribute("weight")[:size]
            columns.append((weight.view(torch.int16).to(torch.int32) & 0xFFFF if weight.dtype == torch.uint16 else weight).reshape(-1, 1).to(torch.float32))
        if "color" in self._names:
            columns.append(self.attribute("color")[:size].reshape(-1, 3).to(torch.float32))
        return torch.cat(columns, 1)
