ssh mp4j@127.0.0.1 "kill -9 10433"
ssh mp4j@127.0.0.1 "kill -9 10434"
kill -9 10423
