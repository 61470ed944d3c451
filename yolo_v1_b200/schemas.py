"""Value types of the inference path: BoundingBox and Detection.

Same fields, constraints and helper methods as the reference's pydantic models
(/root/reference/src/yolo/schemas.py:6-89): coordinates and confidence are
validated to [0, 1] (pydantic raises ValidationError, a ValueError, otherwise --
tests/test_yolo.py:126-138 relies on it), class_id >= 0, equality by value.
"""
from __future__ import annotations

from typing import Annotated, Optional, Tuple

from pydantic import BaseModel, Field

Unit = Annotated[float, Field(ge=0.0, le=1.0)]


class BoundingBox(BaseModel):
    """Axis-aligned box, centre format, normalised to the image (all fields in [0, 1])."""

    x: Unit = Field(..., description="centre x")
    y: Unit = Field(..., description="centre y")
    width: Unit = Field(..., description="box width")
    height: Unit = Field(..., description="box height")

    def to_corners(self) -> Tuple[float, float, float, float]:
        """(x1, y1, x2, y2), still normalised (schemas.py:18-29)."""
        half_w, half_h = self.width / 2, self.height / 2
        return (self.x - half_w, self.y - half_h, self.x + half_w, self.y + half_h)

    def to_pixel_coords(self, img_width: int, img_height: int) -> Tuple[int, int, int, int]:
        """Corners scaled to pixels and truncated with int() (schemas.py:31-50)."""
        x1, y1, x2, y2 = self.to_corners()
        return (int(x1 * img_width), int(y1 * img_height), int(x2 * img_width), int(y2 * img_height))

    @property
    def area(self) -> float:
        return self.width * self.height

    @classmethod
    def from_corners(cls, x1: float, y1: float, x2: float, y2: float) -> "BoundingBox":
        w, h = x2 - x1, y2 - y1
        return cls(x=x1 + w / 2, y=y1 + h / 2, width=w, height=h)

    def __str__(self) -> str:
        return "({:.2f}, {:.2f}, {:.2f}, {:.2f})".format(*self.to_corners())


class Detection(BaseModel):
    """One detected object."""

    class_id: int = Field(..., ge=0)
    class_name: Optional[str] = Field(None)
    confidence: Unit = Field(...)
    bbox: BoundingBox = Field(...)
