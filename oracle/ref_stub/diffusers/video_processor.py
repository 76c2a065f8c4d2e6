from .image_processor import VaeImageProcessor


class VideoProcessor(VaeImageProcessor):
    def postprocess_video(self, video, output_type="np"):
        return video
