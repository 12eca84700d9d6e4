ssh mp4j@127.0.0.1 "kill -9 10465"
ssh mp4j@127.0.0.1 "kill -9 10466"
kill -9 10455
