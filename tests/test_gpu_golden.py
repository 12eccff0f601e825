"""The reference's known-answer scenarios through the real CUDA library (libt4c.so) on a B200."""
import pytest

from tesseract_b200 import capi
from tesseract_b200.managers import B200ContactManager

import scenarios as sc

pytestmark = pytest.mark.gpu


def make(builder, cast):
    return B200ContactManager(builder, cast)


@pytest.mark.parametrize("convex", [False, True])
@pytest.mark.parametrize("tt", [capi.TEST_FIRST, capi.TEST_CLOSEST, capi.TEST_ALL])
def test_box_box(convex, tt):
    sc.box_box(make, convex, tt)


def test_sphere_sphere_primitive():
    sc.sphere_sphere_primitive(make)


def test_sphere_sphere_convex():
    sc.sphere_sphere_convex(make)


@pytest.mark.parametrize("convex", [False, True])
def test_box_sphere(convex):
    sc.box_sphere(make, convex)


@pytest.mark.parametrize("kind", ["cylinder", "cone", "capsule"])
def test_box_cyl_cone_capsule(kind):
    sc.box_cylinder_cone_capsule(make, kind)


def test_mesh_mesh():
    sc.mesh_mesh(make)


@pytest.mark.parametrize("convex,tol", [(False, 0.001), (True, 0.02)])
def test_octomap_sphere(convex, tol):
    sc.octomap_sphere(make, convex, tol)


def test_octomap_mesh_2712():
    sc.octomap_mesh(make)


@pytest.mark.parametrize("cast", [False, True])
def test_octomap_octomap(cast):
    sc.octomap_octomap(make, cast)


@pytest.mark.parametrize("cast", [False, True])
def test_compound_compound(cast):
    sc.compound_compound(make, cast)


@pytest.mark.parametrize("convex", [False, True])
def test_large_dataset_300(convex):
    sc.large_dataset(make, convex)


def test_multi_clone_2700():
    sc.multi_clone(make)


def test_box_box_cast():
    sc.box_box_cast(make)


@pytest.mark.parametrize("convex", [False, True])
def test_sphere_sphere_cast(convex):
    sc.sphere_sphere_cast(make, convex)
