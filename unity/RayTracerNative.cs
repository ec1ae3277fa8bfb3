// RayTracerNative.cs — managed side of the drop-in boundary (include/rtb200.h).  SOURCE ONLY: this image has no
// .NET/Mono/Unity toolchain, so this file has never been compiled here; it is the [DllImport] binding a maintainer of
// vectorized-runner/unity-raytracer adds under Assets/RayTracer/ next to RayTracingSetup.cs (see INTEGRATION.md).
//
// float3 is three packed floats, so Sphere (16 B), Triangle (36 B), PointLightData (24 B), AABB (24 B), CameraData (48 B)
// and UnityEngine.Color (16 B) are blittable and are passed as pinned arrays without `unsafe` (RayTracer.asmdef forbids
// it).  MaterialData holds a `bool` (MaterialData.cs:14) and is NOT blittable: NativeMaterial mirrors it with an int.
using System;
using System.Collections.Generic;
using System.Runtime.InteropServices;
using Unity.Mathematics;
using UnityEngine;

namespace RayTracer
{
	[StructLayout(LayoutKind.Sequential)]
	public struct NativeMaterial                     // rt_material, 56 bytes
	{
		public float3 Diffuse, Ambient, Mirror, Specular;
		public float PhongExponent;
		public int IsMirror;

		public NativeMaterial(MaterialData m)
		{
			Diffuse = m.DiffuseReflectance; Ambient = m.AmbientReflectance; Mirror = m.MirrorReflectance;
			Specular = m.SpecularReflectance; PhongExponent = m.PhongExponent; IsMirror = m.IsMirror ? 1 : 0;
		}
	}

	[StructLayout(LayoutKind.Sequential)]
	public struct NativeRenderParams                 // rt_render_params, 88 bytes
	{
		public float3 Position, Forward, Right, Up;  // CameraData.cs:7-10
		public int ResX, ResY;                       // Resolution.cs:8-9
		public float DistanceToCamera, HalfHorizontalLength, HalfVerticalLength;   // ImagePlane.cs:13-21
		public float3 BackgroundRgb;                 // RayTracingSetup.BackgroundColor (0..1)
		public int MaxReflectionBounces;             // RayTracingSetup.cs:23
		public int SppGrid;                          // 1 = reference behaviour
	}

	public static class RayTracerNative
	{
		const string Lib = "rtb200";                 // librtb200.so / rtb200.dll in Assets/Plugins/x86_64

		[DllImport(Lib)] static extern int rt_create(out IntPtr ctx, int[] deviceIds, int nDevices);
		[DllImport(Lib)] static extern void rt_destroy(IntPtr ctx);
		[DllImport(Lib)] static extern IntPtr rt_last_error(IntPtr ctx);
		[DllImport(Lib)] static extern int rt_scene_begin(IntPtr ctx);
		[DllImport(Lib)] static extern int rt_scene_spheres(IntPtr ctx, Sphere[] centerR2, NativeMaterial[] materials, int n);
		[DllImport(Lib)] static extern int rt_scene_triangles(IntPtr ctx, Triangle[] v, float3[] normals, NativeMaterial[] materials, int n);
		[DllImport(Lib)] static extern int rt_scene_mesh(IntPtr ctx, Triangle[] v, float3[] normals, int nTris, ref NativeMaterial material, ref AABB aabb);
		[DllImport(Lib)] static extern int rt_scene_base_mesh(IntPtr ctx, Triangle[] objectSpaceTriangles, int nTris);
		[DllImport(Lib)] static extern int rt_scene_mesh_instance(IntPtr ctx, int baseMesh, float[] m3x4, ref NativeMaterial material, IntPtr aabbOrNull);
		[DllImport(Lib)] static extern int rt_scene_update_instance(IntPtr ctx, int meshIndex, float[] m3x4, IntPtr aabbOrNull);
		[DllImport(Lib)] static extern int rt_scene_update_instances(IntPtr ctx, int[] meshIndices, float[] m3x4, IntPtr aabbsOrNull, int n);
		[DllImport(Lib)] static extern int rt_scene_lights(IntPtr ctx, PointLightData[] posIntensity, int n, ref float3 ambientRadiance);
		[DllImport(Lib)] static extern int rt_scene_commit(IntPtr ctx, ref AABB sceneAabb);
		[DllImport(Lib)] static extern int rt_scene_update_mesh(IntPtr ctx, int meshIndex, Triangle[] v, float3[] normals, ref AABB aabb);
		[DllImport(Lib)] static extern int rt_scene_update_triangles(IntPtr ctx, int first, int n, Triangle[] v, float3[] normals);
		[DllImport(Lib)] static extern int rt_scene_update_spheres(IntPtr ctx, int first, int n, Sphere[] centerR2);
		[DllImport(Lib)] static extern int rt_scene_update_lights(IntPtr ctx, PointLightData[] posIntensity, int n, ref float3 ambientRadiance);
		[DllImport(Lib)] static extern int rt_scene_recommit(IntPtr ctx, ref AABB sceneAabb);
		[DllImport(Lib)] static extern int rt_scene_refit(IntPtr ctx, ref AABB sceneAabb);
		[DllImport(Lib)] static extern int rt_host_register(IntPtr ctx, IntPtr ptr, UIntPtr bytes);
		[DllImport(Lib)] static extern int rt_host_unregister(IntPtr ctx, IntPtr ptr);
		[DllImport(Lib)] static extern int rt_render(IntPtr ctx, ref NativeRenderParams p, Color[] outRgba32f, IntPtr outRgba8,
		                                             IntPtr outHitId, IntPtr outHitT, IntPtr stats);

		static IntPtr ctx = IntPtr.Zero;

		static void Check(int rc)
		{
			if (rc != 0) throw new InvalidOperationException("librtb200: " + Marshal.PtrToStringAnsi(rt_last_error(ctx)));
		}

		// Scratch arrays that only ever grow: the reference refills its List<T>s every frame (RayTracingSetup.cs:159-169,258-268), and
		// List<T>.ToArray() would allocate a fresh array per list per frame.  List<T>.CopyTo fills the scratch array in place; the
		// native side reads only the first `n` entries, so the arrays may be longer than the lists.
		static Sphere[] sphereBuf = new Sphere[0];
		static Triangle[] triBuf = new Triangle[0];
		static float3[] normalBuf = new float3[0];
		static PointLightData[] lightBuf = new PointLightData[0];
		static NativeMaterial[] sphereMatBuf = new NativeMaterial[0], triMatBuf = new NativeMaterial[0];

		static T[] Fill<T>(List<T> src, ref T[] buf)
		{
			if (buf.Length < src.Count) buf = new T[Math.Max(src.Count, 2 * buf.Length)];
			src.CopyTo(buf);
			return buf;
		}

		static NativeMaterial[] Convert(List<MaterialData> mats, ref NativeMaterial[] buf)
		{
			if (buf.Length < mats.Count) buf = new NativeMaterial[Math.Max(mats.Count, 2 * buf.Length)];
			for (int i = 0; i < mats.Count; i++) buf[i] = new NativeMaterial(mats[i]);
			return buf;
		}

		// rows of the 3x4 part of a Matrix4x4 (UnityEngine stores it column-major: m00 m10 m20 m30 m01 ...)
		static float[] Rows3x4(Matrix4x4 m)
		{
			return new[] { m.m00, m.m01, m.m02, m.m03, m.m10, m.m11, m.m12, m.m13, m.m20, m.m21, m.m22, m.m23 };
		}

		/// Replaces what UpdateScene() hands to the CPU path (RayTracingSetup.cs:120-128): call after Scene.CalculateAABB().
		public static void UploadScene(ref Scene scene)
		{
			if (ctx == IntPtr.Zero) Check(rt_create(out ctx, new[] { 0 }, 1));
			Check(rt_scene_begin(ctx));
			Check(rt_scene_spheres(ctx, Fill(scene.SphereData.Spheres, ref sphereBuf), Convert(scene.SphereData.Materials, ref sphereMatBuf), scene.SphereData.Spheres.Count));
			Check(rt_scene_triangles(ctx, Fill(scene.TriangleData.Triangles, ref triBuf), Fill(scene.TriangleData.Normals, ref normalBuf),
			                         Convert(scene.TriangleData.Materials, ref triMatBuf), scene.TriangleData.Triangles.Count));
			foreach (var mesh in scene.MeshData.Meshes)          // list order == ObjectId.MeshIndex (Scene.cs:64)
			{
				var mat = new NativeMaterial(mesh.MaterialData);
				var box = mesh.AABB;
				Check(rt_scene_mesh(ctx, mesh.Triangles, mesh.TriangleNormals, mesh.Triangles.Length, ref mat, ref box));
			}
			var ambient = scene.AmbientLight.Radiance;
			Check(rt_scene_lights(ctx, Fill(scene.PointLights, ref lightBuf), scene.PointLights.Count, ref ambient));
			var aabb = scene.AABB;                               // Scene.cs:15, computed by Scene.CalculateAABB()
			Check(rt_scene_commit(ctx, ref aabb));
		}

		// ---- instanced meshes: SceneMesh.Mesh (SceneMesh.cs:11-53) without the per-frame world bake (INTEGRATION.md section 6) ----
		// Between rt_scene_begin and rt_scene_commit, in place of the rt_scene_mesh call of that SceneMesh:
		//   int b = BaseMesh(objectSpaceTriangles)            once per distinct MeshFilter.sharedMesh (gathered through its index buffer, SceneMesh.cs:36-41)
		//   MeshInstance(b, transform.localToWorldMatrix, material)      once per SceneMesh, in MeshIndex order (baked meshes and instances share it)
		// Later frames: UpdateInstance(meshIndex, transform.localToWorldMatrix) for the ones that moved, then Recommit().
		public static int BaseMesh(Triangle[] objectSpaceTriangles)
		{
			int rc = rt_scene_base_mesh(ctx, objectSpaceTriangles, objectSpaceTriangles.Length);
			if (rc < 0) Check(rc);
			return rc;
		}

		/// Mesh.AABB is computed by the library exactly as SceneMesh.cs:22-31 does (min / max of the baked vertices).
		public static void MeshInstance(int baseMesh, Matrix4x4 localToWorld, MaterialData material)
		{
			var mat = new NativeMaterial(material);
			Check(rt_scene_mesh_instance(ctx, baseMesh, Rows3x4(localToWorld), ref mat, IntPtr.Zero));
		}

		public static void UpdateInstance(int meshIndex, Matrix4x4 localToWorld)
		{
			Check(rt_scene_update_instance(ctx, meshIndex, Rows3x4(localToWorld), IntPtr.Zero));
		}

		/// Every moved instance in ONE native call (rt_scene_update_instances; all or nothing).  The arrays are grow-only scratch: no
		/// allocation per frame once they have reached the scene's instance count.
		static int[] instIdx = new int[0];
		static float[] instRows = new float[0];
		public static void UpdateInstances(IList<int> meshIndices, IList<Matrix4x4> localToWorld)
		{
			int n = meshIndices.Count;
			if (instIdx.Length < n) { instIdx = new int[n]; instRows = new float[12 * n]; }
			for (int i = 0; i < n; i++)
			{
				instIdx[i] = meshIndices[i];
				Matrix4x4 m = localToWorld[i];
				for (int r = 0; r < 3; r++)
					for (int c = 0; c < 4; c++)
						instRows[12 * i + 4 * r + c] = m[r, c];
			}
			Check(rt_scene_update_instances(ctx, instIdx, instRows, IntPtr.Zero, n));
		}

		/// After UpdateInstance calls: rebuilds the instance tree (microseconds) and whatever else was updated.  sceneAabb = Scene.AABB.
		public static void Recommit(AABB sceneAabb)
		{
			Check(rt_scene_recommit(ctx, ref sceneAabb));
		}

		/// Dirty-tracking alternative to UploadScene for frames in which the object COUNTS did not change (the usual case: only
		/// transforms moved).  meshMoved[i] = SceneMesh i's transform.hasChanged; spheres / loose triangles / lights are small
		/// and are simply re-sent.  The BVH is rebuilt on the device from the resident arrays (0.7 ms per million primitives).
		/// refitOnly: keep the tree and only recompute its boxes (rt_scene_refit) — for small per-frame motion; pass false
		/// every so often (or after large motion) so that the tree is rebuilt for the new layout.
		public static void UpdateScene(ref Scene scene, bool[] meshMoved, bool refitOnly = false)
		{
			var meshes = scene.MeshData.Meshes;
			for (int i = 0; i < meshes.Count; i++)
			{
				if (!meshMoved[i]) continue;
				var box = meshes[i].AABB;
				Check(rt_scene_update_mesh(ctx, i, meshes[i].Triangles, meshes[i].TriangleNormals, ref box));
			}
			Check(rt_scene_update_spheres(ctx, 0, scene.SphereData.Spheres.Count, Fill(scene.SphereData.Spheres, ref sphereBuf)));
			Check(rt_scene_update_triangles(ctx, 0, scene.TriangleData.Triangles.Count, Fill(scene.TriangleData.Triangles, ref triBuf),
			                                Fill(scene.TriangleData.Normals, ref normalBuf)));
			var ambient = scene.AmbientLight.Radiance;
			Check(rt_scene_update_lights(ctx, Fill(scene.PointLights, ref lightBuf), scene.PointLights.Count, ref ambient));
			var aabb = scene.AABB;
			if (refitOnly) Check(rt_scene_refit(ctx, ref aabb));
			else Check(rt_scene_recommit(ctx, ref aabb));
		}

		/// Replaces CastPixelRays(CameraData) (RayTracingSetup.cs:275-302): fills PixelColors (x + y * ResX, y = 0 top row).
		public static void CastPixelRays(CameraData cam, ImagePlane plane, Color background, int maxBounces, Color[] pixelColors)
		{
			var p = new NativeRenderParams
			{
				Position = cam.Position, Forward = cam.Forward, Right = cam.Right, Up = cam.Up,
				ResX = plane.Resolution.X, ResY = plane.Resolution.Y,
				DistanceToCamera = plane.DistanceToCamera, HalfHorizontalLength = plane.HalfHorizontalLength,
				HalfVerticalLength = plane.HalfVerticalLength,
				BackgroundRgb = new float3(background.r, background.g, background.b),
				MaxReflectionBounces = maxBounces, SppGrid = 1
			};
			Check(rt_render(ctx, ref p, pixelColors, IntPtr.Zero, IntPtr.Zero, IntPtr.Zero, IntPtr.Zero));
		}

		static GCHandle pinnedPixels;   // PixelColors is allocated once per resolution (RayTracingSetup.cs:206-210): pin + page-lock it once

		/// Optional: call after PixelColors is (re)allocated.  The frame copy then runs at PCIe speed (3.1 instead of 6.5 ms at 4K).
		public static void RegisterPixelColors(Color[] pixelColors)
		{
			if (ctx == IntPtr.Zero) Check(rt_create(out ctx, new[] { 0 }, 1));
			if (pinnedPixels.IsAllocated) { rt_host_unregister(ctx, pinnedPixels.AddrOfPinnedObject()); pinnedPixels.Free(); }
			pinnedPixels = GCHandle.Alloc(pixelColors, GCHandleType.Pinned);
			Check(rt_host_register(ctx, pinnedPixels.AddrOfPinnedObject(), (UIntPtr)((ulong)pixelColors.Length * 16UL)));
		}

		public static void Shutdown()
		{
			if (pinnedPixels.IsAllocated) { rt_host_unregister(ctx, pinnedPixels.AddrOfPinnedObject()); pinnedPixels.Free(); }
			if (ctx != IntPtr.Zero) { rt_destroy(ctx); ctx = IntPtr.Zero; }
		}
	}
}
