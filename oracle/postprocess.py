"""ORACLE (test infrastructure only) -- restatement of yolox.utils.postprocess.

Upstream file yolox/utils/boxes.py::postprocess (Megvii YOLOX, un-vendored; call
sites /root/reference/src/yolox_detector.py:309-314,381-386).  NMS itself is the real
torchvision.ops implementation present in this image.  PARITY UNPINNED (no reference
golden vectors).

On a GPU the reference's batched_nms takes torchvision's coordinate-offset branch
(numel <= 100000, torchvision/ops/boxes.py:80-83); on CPU with > 4000 elements it
takes the per-class loop.  The CUDA build follows the GPU branch, so this oracle
forces the coordinate trick regardless of device.
"""
import torch
import torchvision


def batched_nms_coordinate_trick(boxes, scores, idxs, iou_threshold):
    if boxes.numel() == 0:
        return torch.empty((0,), dtype=torch.int64, device=boxes.device)
    max_coordinate = boxes.max()
    offsets = idxs.to(boxes) * (max_coordinate + torch.tensor(1).to(boxes))
    boxes_for_nms = boxes + offsets[:, None]
    return torchvision.ops.nms(boxes_for_nms, scores, iou_threshold)


def postprocess(prediction, num_classes, conf_thre=0.7, nms_thre=0.45, class_agnostic=False):
    prediction = prediction.clone()
    box_corner = prediction.new(prediction.shape)
    box_corner[:, :, 0] = prediction[:, :, 0] - prediction[:, :, 2] / 2
    box_corner[:, :, 1] = prediction[:, :, 1] - prediction[:, :, 3] / 2
    box_corner[:, :, 2] = prediction[:, :, 0] + prediction[:, :, 2] / 2
    box_corner[:, :, 3] = prediction[:, :, 1] + prediction[:, :, 3] / 2
    prediction[:, :, :4] = box_corner[:, :, :4]

    output = [None for _ in range(len(prediction))]
    for i, image_pred in enumerate(prediction):
        if not image_pred.size(0):
            continue
        class_conf, class_pred = torch.max(image_pred[:, 5:5 + num_classes], 1, keepdim=True)
        conf_mask = (image_pred[:, 4] * class_conf.squeeze() >= conf_thre).squeeze()
        detections = torch.cat((image_pred[:, :5], class_conf, class_pred.float()), 1)
        detections = detections[conf_mask]
        if not detections.size(0):
            continue
        if class_agnostic:
            keep = torchvision.ops.nms(detections[:, :4], detections[:, 4] * detections[:, 5], nms_thre)
        else:
            keep = batched_nms_coordinate_trick(
                detections[:, :4], detections[:, 4] * detections[:, 5], detections[:, 6], nms_thre)
        detections = detections[keep]
        output[i] = detections if output[i] is None else torch.cat((output[i], detections))
    return output
